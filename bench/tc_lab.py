"""Map which shared-memory word tcgen05.mma reads for every (row, k) of an operand under a given descriptor.

One operand image is a ramp (value = word index, split into two exact TF32 runs), the other an identity, so
D is the operand as the tensor core sees it. Prints, per experiment, the decoded word index for a few rows
and the fraction of entries that match the hypothesised address function.
"""
import json
import sys

import torch

sys.path.insert(0, ".")
from distributed_neural_network_b200.ops import native  # noqa: E402

A_WORDS = 8192


def desc(layout_type, lbo, sbo):
    return ((lbo >> 4) & 0x3FFF) << 16 | ((sbo >> 4) & 0x3FFF) << 32 | 1 << 46 | layout_type << 61


def idesc(M, N, a_mn=0, b_mn=0):
    return (1 << 4) | (2 << 7) | (2 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def identity_kmajor(rows):
    """[rows][8] identity (row == k), canonical no-swizzle K-major; returns image, LBO, SBO."""
    groups = rows // 8
    img = torch.zeros(2 * groups * 32)
    for k in range(8):
        r = k
        img[((k >> 2) * groups + (r >> 3)) * 32 + (r & 7) * 4 + (k & 3)] = 1.0
    return img, groups * 128, 128


def ramp_runs():
    i = torch.arange(A_WORDS)
    return (i & 1023).float(), (i >> 10).float()


def decode_a(layout_type, lbo, sbo, a_mn):
    """Word index read for A(m, k), m < 128, k < 8."""
    bimg, blbo, bsbo = identity_kmajor(16)
    lo, hi = ramp_runs()
    out = []
    for img in (lo, hi):
        D = native.tc_lab(img.cuda(), bimg.cuda(), desc(layout_type, lbo, sbo), desc(0, blbo, bsbo),
                          idesc(128, 16, a_mn, 0), 16)
        torch.cuda.synchronize()
        out.append(D[:, :8].cpu())
    return (out[1] * 1024 + out[0]).long()          # [128][8]


def decode_b(layout_type, lbo, sbo, b_mn, N=128):
    """Word index read for B(n, k), n < N, k < 8."""
    aimg, albo, asbo = identity_kmajor(128)
    lo, hi = ramp_runs()
    out = []
    for img in (lo, hi):
        D = native.tc_lab(aimg.cuda(), img.cuda(), desc(0, albo, asbo), desc(layout_type, lbo, sbo),
                          idesc(128, N, 0, b_mn), N)
        torch.cuda.synchronize()
        out.append(D[:8, :].cpu().t())
    return (out[1] * 1024 + out[0]).long()          # [N][8]


def hyp_mn_sw128_32b(rows, lbo, sbo):
    m = torch.arange(rows).view(-1, 1)
    k = torch.arange(8).view(1, -1)
    byte = (m // 32) * lbo + (k // 4) * sbo + (k % 4) * 128 + ((((m % 32) // 8) ^ (k % 4)) * 32) + (m % 8) * 4
    return byte // 4


def hyp_k_nosw(rows, lbo, sbo):
    m = torch.arange(rows).view(-1, 1)
    k = torch.arange(8).view(1, -1)
    return ((k // 4) * lbo + (m // 8) * sbo + (m % 8) * 16 + (k % 4) * 4) // 4


def show(name, got, hyp=None):
    rows = [0, 1, 4, 7, 8, 9, 16, 31, 32, 33, 64, 127]
    rec = {"name": name, "all_zero": bool((got == 0).all()),
           "rows": {str(r): got[r].tolist() for r in rows if r < got.shape[0]}}
    if hyp is not None:
        rec["match"] = float((got == hyp).float().mean())
    print(json.dumps(rec))


def main():
    show("A K-major nosw (sanity)", decode_a(0, 2048, 128, 0), hyp_k_nosw(128, 2048, 128))
    show("A MN sw128_base32 lbo4096 sbo512", decode_a(1, 4096, 512, 1), hyp_mn_sw128_32b(128, 4096, 512))
    show("A MN sw128_base32 lbo512 sbo4096", decode_a(1, 512, 4096, 1), hyp_mn_sw128_32b(128, 512, 4096))
    show("B MN sw128_base32 lbo4096 sbo512 N128", decode_b(1, 4096, 512, 1), hyp_mn_sw128_32b(128, 4096, 512))
    show("B MN sw128_base32 lbo4096 sbo512 N32", decode_b(1, 4096, 512, 1, 32), hyp_mn_sw128_32b(32, 4096, 512))
    show("A MN sw128 lbo4096 sbo1024", decode_a(2, 4096, 1024, 1))
    show("A MN nosw lbo128 sbo2048", decode_a(0, 128, 2048, 1))
    show("A MN nosw lbo2048 sbo128", decode_a(0, 2048, 128, 1))


if __name__ == "__main__":
    main()
