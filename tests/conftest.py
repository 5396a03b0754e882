import os
import sys

# The threaded virtual-rank tests run up to 8 handles (16 streams) whose persistent kernels wait for each other on ONE
# device: give every stream its own hardware queue, or kernels of different "ranks" that share a queue serialise and
# deadlock (must be set before the CUDA context exists; irrelevant to the one-process-per-GPU production layout).
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
# ... and load every kernel when its module is loaded, not at its first launch: a first launch that has to wait for the device
# to drain while another "rank" of the same process spins on a flag that launch would raise never returns (libfeab also
# touches its stepping kernels at handle creation, which is what protects a process that initialised CUDA before this line).
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
