"""One rank of the reference's own distributed test file, UNMODIFIED (tests/distributed/tests_fft.py), running on
the product's process-group registry (makani_b200.comm as makani.utils.comm) and the product's
``distributed_transpose_polar / _azimuth`` (as torch_harmonics.distributed, imported by makani/mpu/layers.py:34-35 and
makani/mpu/fft3d.py:8-9).  Launched by tests/test_reference_dropin.py with torchrun (gloo, CPU):

    GRID_H=2 GRID_W=1 python -m torch.distributed.run --nproc-per-node 2 ... run_reference_tests_fft.py [pattern ...]

The file pins the transposes' forward and backward through distributed 2-D / 3-D FFTs against the local transforms at
rel-L2 <= 1e-6.  Exit status 0 = every selected test passed on this rank."""
import os
import sys
import unittest
import warnings

warnings.filterwarnings("ignore")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import refharness as H                                                     # noqa: E402

H.install("product")
import torch                                                                # noqa: E402
import torch.distributed as dist                                            # noqa: E402

# The file's own gather helpers call dist.all_gather with per-rank shapes (361 = 181 + 180).  NCCL accepts that, gloo
# (the only backend without GPUs) does not - so under gloo, and only for uneven shapes, all_gather is emulated with one
# broadcast per rank.  This touches the test's bookkeeping, not the transposes under test.
_all_gather = dist.all_gather


def _all_gather_uneven_ok(tensor_list, tensor, group=None, async_op=False):
    if dist.get_backend(group) != "gloo" or all(t.shape == tensor.shape for t in tensor_list):
        return _all_gather(tensor_list, tensor, group=group, async_op=async_op)
    me = dist.get_rank(group=group)
    for i, t in enumerate(tensor_list):
        buf = tensor.contiguous().clone() if i == me else torch.empty_like(t)
        real = torch.view_as_real(buf) if buf.is_complex() else buf
        dist.broadcast(real, src=dist.get_global_rank(group, i) if group is not None else i, group=group)
        if t is not tensor:
            t.copy_(buf)
    return None


dist.all_gather = _all_gather_uneven_ok
mod = H.load_file("tests/distributed/tests_fft.py", "reference_tests_fft")
assert mod.__file__.startswith("/root/reference/")
patterns = sys.argv[1:]
loader = unittest.TestLoader()
names = [n for n in loader.getTestCaseNames(mod.TestDistributedRealFFT)
         if not patterns or any(p in n for p in patterns)]
assert names, "no test selected"
suite = unittest.TestSuite(mod.TestDistributedRealFFT(n) for n in names)
rank = int(os.environ.get("RANK", "0"))
stream = sys.stderr if rank == 0 else open(os.devnull, "w")
res = unittest.TextTestRunner(stream=stream, verbosity=2).run(suite)
if dist.is_initialized():
    dist.barrier()
    dist.destroy_process_group()
ok = res.wasSuccessful() and res.testsRun == len(names)
print(f"rank {rank}: ran {res.testsRun} reference tests ({', '.join(names)}): {'OK' if ok else 'FAILED'}")
sys.exit(0 if ok else 1)
