"""-m gpu: hn_op_encode_noalloc (csrc/hn_noalloc.cu) - the hash-grid encoding through the library's static __device__ scratch, for
callers that cannot allocate device memory. No counterpart in the reference.

Checked through the C ABI against the existing entry point hn_op_encode, which is itself bit-exact against the oracle and the reference's
kernel_grid (tests/test_gpu_parity.py, tests/test_gpu_refgpu.py): the encoding is fp16 integer-like work, so the tolerance is the one those
tests use - bit-exact - on several calls in a row (the calls share one scratch array that keeps the previous call's contents).
This file sorts last on purpose (a failure here must not hide other tests); it builds everything it needs
itself, so it passes alone in a fresh process and after the rest of the suite."""
import ctypes
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_device_bytes():
    """free bytes of device 0 (cudaMemGetInfo through the CUDA runtime: a device-wide figure, it sees the library's allocations)"""
    rt = None
    for name in ("libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so"):
        try:
            rt = ctypes.CDLL(name)
            break
        except OSError:
            continue
    if rt is None:
        import torch
        return torch.cuda.mem_get_info(0)[0]
    free, total = ctypes.c_size_t(0), ctypes.c_size_t(0)
    assert rt.cudaMemGetInfo(ctypes.byref(free), ctypes.byref(total)) == 0
    return free.value


@pytest.mark.parametrize("config", ["nerf/base.json", "geometry/base.json"])
def test_encode_noalloc_matches_encode_on_calls_in_a_row(hn, root, config):
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", *config.split("/")))
    info = tb.network_info()
    tb.set_params(hn.synthetic.random_params(info, 1337, 1.0))
    rng = np.random.default_rng(11)
    n_max = hn.NOALLOC_MAX_POSITIONS
    # call 1 (warm-up: the kernel's module is loaded lazily on its first launch)
    p1 = rng.random((50_000, 3), dtype=np.float32)
    a1 = tb.op_encode_noalloc(p1)
    _free_device_bytes()  # the reader's own first-use set-up is not part of the measurement
    free_before = _free_device_bytes()
    # call 2: the largest call the scratch takes; call 3: a small one right after it (stale rows of call 2 stay in the array behind it)
    p2 = rng.random((n_max, 3), dtype=np.float32)
    a2 = tb.op_encode_noalloc(p2)
    p3 = np.concatenate([rng.random((999, 3), dtype=np.float32), np.array([[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.5]], np.float32)])
    a3 = tb.op_encode_noalloc(p3)
    a1_again = tb.op_encode_noalloc(p1)
    free_after = _free_device_bytes()
    assert free_after >= free_before, f"device memory was allocated by the no-alloc entry point: {free_before - free_after} bytes"
    # the existing entry point (grows the context's staging buffers with cudaMalloc) on the same inputs: bit-exact
    for name, p, a in (("call 1", p1, a1), ("call 2", p2, a2), ("call 3", p3, a3), ("call 1 again", p1, a1_again)):
        want = tb.op_encode(p)
        assert a.shape == want.shape == (p.shape[0], 32)
        assert np.array_equal(a.view(np.uint16), want.view(np.uint16)), f"{name}: hn_op_encode_noalloc differs from hn_op_encode"
    assert np.abs(a2.astype(np.float32)).max() > 0.05  # not a trivially zero encoding
    # n = 0 is a no-op, sizes beyond the static scratch are rejected with a clear error (and the context stays usable)
    assert tb.op_encode_noalloc(np.zeros((0, 3), np.float32)).shape == (0, 32)
    with pytest.raises(RuntimeError, match="HN_NOALLOC_MAX_POSITIONS"):
        tb.op_encode_noalloc(np.zeros((n_max + 1, 3), np.float32))
    assert np.array_equal(tb.op_encode_noalloc(p3).view(np.uint16), a3.view(np.uint16))


def test_encode_noalloc_needs_parameters(hn, root):
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    with pytest.raises(RuntimeError, match="parameters"):
        tb.op_encode_noalloc(np.zeros((4, 3), np.float32))
