"""Worker for the NVLS test: all-reduce the same random vectors with the one-shot and the NVLS variant."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402

from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402


def main():
    out_dir = sys.argv[1]
    os.makedirs(out_dir, exist_ok=True)   # every rank writes its result here; nobody may rely on another rank's mkdir
    tr = Trainer(TrainConfig(device="cuda", synthetic=True, quiet=True, sync_mode="step-allreduce", seed=3,
                             peer_timeout_s=20.0, log_dir=out_dir + "/log"))
    from distributed_neural_network_b200.parallel import nvls
    res = {"rank": tr.rank, "nvls": tr.nvls_ok, "why_not": nvls.last_error}
    if not tr.nvls_ok:
        print(f"[rank {tr.rank}] NVLS group not created: {nvls.last_error or 'another rank failed'}", flush=True)
    if tr.nvls_ok:
        eng = tr.engine
        g = torch.Generator().manual_seed(100 + tr.rank)
        outs = {}
        for n in (L.FLAT_NUMEL, 4096, 512 + 64):
            for rep in range(3):                      # several rounds: both parities, flag reuse
                x = torch.randn(L.FLAT_NUMEL, generator=g).cuda()
                y_one = torch.zeros(L.FLAT_NUMEL, device="cuda")
                y_nvls = torch.zeros(L.FLAT_NUMEL, device="cuda")
                eng.allreduce(x[:n], y_one[:n], mode=0, two_shot=0)
                eng.allreduce(x[:n], y_nvls[:n], mode=0, two_shot="nvls")
                torch.cuda.synchronize()
                outs[f"{n}_{rep}"] = (x.cpu(), y_one.cpu(), y_nvls.cpu())
        res["outs"] = outs
        res["status"] = int(eng.read_ctl().status)
    torch.save(res, f"{out_dir}/rank{tr.rank}.pt")
    tr.close()
    tr.comm.close()


if __name__ == "__main__":
    main()
