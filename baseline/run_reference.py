"""Run ONE rank of the UNMODIFIED reference script (baseline/_ref) for the benchmark's reference arm.

What is substituted is only what the image lacks: mpi4py (-> baseline/shims/mpi4py over local
sockets), neptune (-> no-op), and the CIFAR-10 download (-> synthetic uint8 images of the same
shape served through the reference's own transform and DataLoader). The script's code, model,
optimizer, loop and exchange run as written (reference: data_parallelism_train.py:56-272), on CPU,
as the reference's distributed path always does.

argv: <script> <out_json> <train_len> <test_len> -- <reference argv...>
"""
import json
import os
import runpy
import sys
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE / "shims"))


def main():
    script, out_json, train_len, test_len = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
    ref_argv = sys.argv[6:] if len(sys.argv) > 5 and sys.argv[5] == "--" else sys.argv[5:]
    import numpy as np
    import torch
    import torchvision

    threads = os.environ.get("REFSHIM_THREADS")
    if threads:
        torch.set_num_threads(int(threads))

    class SyntheticCIFAR10(torch.utils.data.Dataset):
        """CIFAR-10-shaped uint8 HWC images + int labels; honours ``transform`` like torchvision's class."""

        def __init__(self, root=None, train=True, download=False, transform=None, target_transform=None):
            n = train_len if train else test_len
            rng = np.random.RandomState(0 if train else 1)
            self.data = rng.randint(0, 256, (n, 32, 32, 3), dtype=np.uint8)
            self.targets = rng.randint(0, 10, (n,)).tolist()
            self.transform = transform

        def __len__(self):
            return len(self.targets)

        def __getitem__(self, i):
            img = self.data[i]
            if self.transform is not None:
                img = self.transform(img)
            return img, self.targets[i]

    torchvision.datasets.CIFAR10 = SyntheticCIFAR10
    os.makedirs("log", exist_ok=True)
    ref_dir = str(HERE / "_ref")
    sys.path.insert(0, ref_dir)
    sys.argv = [script] + ref_argv
    t0 = time.time()
    runpy.run_path(os.path.join(ref_dir, script), run_name="__main__")
    import mpi4py  # the shim: collect its timestamps

    Path(out_json).write_text(json.dumps({"rank": int(os.environ.get("REFSHIM_RANK", "0")), "t0": t0,
                                          "t1": time.time(), "events": mpi4py._EVENTS}))


if __name__ == "__main__":
    main()
