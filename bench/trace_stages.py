"""Per-stage timeline of the clustered forward/backward kernel (SM clocks, thread 0 of each CTA)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200 import Network  # noqa: E402
from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.ops import native  # noqa: E402

NAMES = ["start->arrive", "cluster_wait", "img+weights", "conv1", "wait p1", "conv2", "wait p2", "fc1(+wait)",
         "fc2(+wait)", "fc3,CE,dx3,dx2(+wait)", "dx1(+wait)", "conv2 wgrad", "dgrad2", "conv1 wgrad"]


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    cs = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    prec = sys.argv[3] if len(sys.argv) > 3 else native.default_precision()
    torch.manual_seed(0)
    m = Network().cuda()
    eng = native.Engine(m.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, B, 0.001, 0.9, precision=prec)
    eng.set_cluster_size(cs)
    x = torch.randn(B, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (B,), device="cuda")
    for _ in range(5):
        eng.step(x, y)
    n_cta = B * cs
    tr = torch.zeros(n_cta * 32, dtype=torch.int64, device="cuda")
    eng.set_trace(tr)
    # four steps chained inside one engine call: the stamps of the LAST one stay in the buffer (a chained step waits on
    # the update counter and, in tf32 mode, fetches its B operands ready-made -- the path a training run takes)
    eng.train_device(x.unsqueeze(0).repeat(4, 1, 1, 1, 1), y.unsqueeze(0).repeat(4, 1))
    torch.cuda.synchronize()
    eng.set_trace(None)
    t = tr.view(n_cta, 32).cpu()
    d = (t[:, 1:14] - t[:, 0:13]).float()
    print(f"B={B} cs={cs} {prec}: per-stage cycles (mean / max over {n_cta} CTAs); total {(t[:,13]-t[:,0]).float().mean():.0f} cycles")
    for k in range(13):
        print(f"  {NAMES[k+1]:28s} {d[:,k].mean():8.0f} {d[:,k].max():8.0f}")
    ent, ext = t[:, 14].double(), t[:, 15].double()
    print(f"  globaltimer: first CTA entry -> last CTA entry {ent.max()-ent.min():.0f} ns; "
          f"entry -> end of image loop, per CTA mean {(ext-ent).mean():.0f} ns; first entry -> last exit {ext.max()-ent.min():.0f} ns")
    pr = t[:, 16:21].double()
    names = ["entry->mbar init+sync", "cluster arrive", "issue image TMA + weight loads issued", "weights landed (sync)"]
    for k in range(4):
        print(f"    prologue {names[k]:40s} {(pr[:, k+1]-pr[:, k]).mean():8.0f} cycles")
    print(f"    prologue w2d build + split consts -> trace0 {(t[:,0]-t[:,20]).double().mean():8.0f} cycles")
    fc = t[:, 22:30].double()
    print("    fc1 warp0: row0 dot %.0f, sum+bias %.0f, st.async %.0f | gap %.0f | row1 dot %.0f, sum %.0f, st.async %.0f (cycles, mean)" % (
        (fc[:,1]-fc[:,0]).mean(), (fc[:,2]-fc[:,1]).mean(), (fc[:,3]-fc[:,2]).mean(), (fc[:,4]-fc[:,3]).mean(),
        (fc[:,5]-fc[:,4]).mean(), (fc[:,6]-fc[:,5]).mean(), (fc[:,7]-fc[:,6]).mean()))
    print(f"    fc1 stage start (trace6) -> first row start {(fc[:,0]-t[:,6].double()).mean():.0f} cycles")
    if prec == "tf32":   # thread 0's MMA issue loops (stamps 30 / 31 / 21 sit just before the commits)
        print(f"    tcgen05: conv1 stage start -> its MMAs issued {(t[:,30]-t[:,2]).double().mean():.0f} cycles; "
              f"conv2: p1 complete (trace4) -> MMAs issued {(t[:,31]-t[:,4]).double().mean():.0f}; "
              f"dgrad: dx1 done (trace10) -> MMAs issued {(t[:,21]-t[:,10]).double().mean():.0f}")
    loop_ns = (t[:, 13] - t[:, 0]).double().mean() / 1.965
    print(f"  image loop {loop_ns:.0f} ns  => prologue (entry -> cluster barrier) ~ {(ext-ent).mean() - loop_ns:.0f} ns")


if __name__ == "__main__":
    main()
