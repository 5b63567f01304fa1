"""How fast are the tiny tcgen05.mma of the fused step (M=128, N=16, K=8, TF32, no-swizzle K-major operands)?

Cycles per MMA for a dependent chain on one accumulator tile vs rotating over several, and for the operand
geometries the step kernel uses (K-chunks 16 bytes apart / a plane apart, row groups 128 / 224 / 528 bytes apart).
"""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from distributed_neural_network_b200.ops import native  # noqa: E402


def desc(layout_type, lbo, sbo):
    return ((lbo >> 4) & 0x3FFF) << 16 | ((sbo >> 4) & 0x3FFF) << 32 | 1 << 46 | layout_type << 61


def idesc(M, N):
    return (1 << 4) | (2 << 7) | (2 << 10) | ((N >> 3) << 17) | ((M >> 4) << 24)


def run(name, lbo, sbo, a_step, nmma, n_acc, N=16, M=128, layout=0):
    lib = native.require()
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    best = None
    for _ in range(5):
        rc = lib.dnnb200_tc_lab_time(64 * 1024, 64 * 1024, C.c_ulonglong(desc(layout, lbo, sbo)),
                                     C.c_ulonglong(desc(0, 128, 25 * 2 * 128)), C.c_uint(idesc(M, N)), nmma, a_step, 256,
                                     n_acc, max(N, 16), C.c_void_p(out.data_ptr()), C.c_void_p(0))
        assert rc == 0, lib.dnnb200_last_error()
        torch.cuda.synchronize()
        v = out.cpu().tolist()
        best = v if best is None or v[1] < best[1] else best
    print(f"{name:58s} nmma {nmma:3d} acc {n_acc}: issue {best[0] / nmma:6.1f}  total {best[1] / nmma:6.1f} cycles/MMA "
          f"({best[1]} cycles)", flush=True)


def main():
    for n_acc in (1, 4):
        run("conv1-like: chunks 16 B apart, row groups 528 B apart", 16, 528, 32, 26, n_acc)
    run("conv1-like but row groups 512 B apart", 16, 512, 32, 26, 1)
    for n_acc in (1, 4):
        run("conv2-like: chunks a plane (4864 B) apart, groups 224 B", 4864, 224, 16, 50, n_acc)
    run("conv2-like but groups 256 B apart", 4864, 256, 16, 50, 1)
    for n_acc in (1, 4):
        run("dgrad-like: chunks a plane (5504 B) apart, groups 128 B", 5504, 128, 16, 50, n_acc)
    run("dgrad-like, operands never move (a_step 0)", 5504, 128, 0, 50, 1)
    for n_acc in (1,):
        run("canonical: chunks 2048 B apart, groups 128 B", 2048, 128, 4096, 12, n_acc)
    run("128B-swizzle K-major (layout 2), groups 1024 B apart", 16, 1024, 32, 12, 1, layout=2)
    for iss in (-1, -2, -4):
        run(f"conv2-like, {-iss} issuing thread(s), own accumulators", 4864, 224, 16, 48, iss)
    # B operand: N rows -> N/8 groups 6400 B apart; keep N <= 64 so that 50 steps of 256 B stay inside the 64 KB image
    for N in (16, 32, 64):
        run(f"conv2-like, N={N}", 4864, 224, 16, 50, 1, N=N)
    for n in (1, 2, 4, 8, 16, 50):
        run("conv2-like chain length", 4864, 224, 16, n, 1)


if __name__ == "__main__":
    main()
