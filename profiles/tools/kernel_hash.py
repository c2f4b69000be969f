"""sha256 over the sources the path kernel is compiled from: ties a committed ncu capture to the code it measured."""
import glob
import hashlib
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(ROOT, "real-time-gpu-ray-tracer_b200", "csrc")


def kernel_source_hash() -> str:
    h = hashlib.sha256()
    for p in sorted(glob.glob(os.path.join(CSRC, "kernels_persist.cu")) + glob.glob(os.path.join(CSRC, "*.h")) + [os.path.join(CSRC, "Makefile")]):
        h.update(os.path.basename(p).encode())
        h.update(open(p, "rb").read())
    return h.hexdigest()


if __name__ == "__main__":
    print(kernel_source_hash())
