"""Worker for the descriptor hand-over test: rank 0 opens a file and passes the descriptor; everybody reads through theirs."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from distributed_neural_network_b200.ops import native  # noqa: E402


def main():
    rank, world, name, path, out_dir = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4], sys.argv[5]
    fd = os.open(path, os.O_RDONLY) if rank == 0 else -1
    word, got = native.share_fd(rank, world, name, word=0xC0FFEE if rank == 0 else 0, fd=fd, timeout_s=20.0)
    data = os.pread(got, 64, 0)          # pread: the duplicates share one file offset
    os.close(got)
    with open(os.path.join(out_dir, f"rank{rank}.txt"), "w") as f:
        f.write(f"{word:x} {got >= 0} {data.decode()}")


if __name__ == "__main__":
    main()
