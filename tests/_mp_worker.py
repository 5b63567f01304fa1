"""Worker for the multi-process CPU tests: trains a few epochs and dumps the final state."""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402

from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402


def main():
    out_dir, mode = sys.argv[1], sys.argv[2]
    os.makedirs(out_dir, exist_ok=True)   # every rank writes its result here; nobody may rely on another rank's mkdir
    kw = json.loads(sys.argv[3]) if len(sys.argv) > 3 else {}
    cfg = TrainConfig(device=kw.pop("device", "cpu"), synthetic=True, quiet=True, sync_mode=mode,
                      log_dir=out_dir + "/log", **kw)
    tr = Trainer(cfg)
    if os.environ.get("DNNB200_PROVIDER_PROBE"):       # which control-plane provider did bootstrap.init() pick?
        Path(os.environ["DNNB200_PROVIDER_PROBE"] + f".{tr.rank}").write_text(tr.comm.provider)
    hist = tr.fit()
    torch.save({"flat": tr.model.flat_params.detach().cpu(), "history": hist, "rank": tr.rank,
                "timers": tr.timers.__dict__, "exchange_table": tr.exchange_table,
                "ll128_ok": tr.ll128_ok, "variant": tr._exchange_variant() if tr.world > 1 else 0}, f"{out_dir}/rank{tr.rank}.pt")
    tr.close()
    tr.comm.close()


if __name__ == "__main__":
    main()
