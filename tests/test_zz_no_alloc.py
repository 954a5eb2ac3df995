"""The no-allocation training step, called straight through the C ABI.

Some users of the library cannot allocate on the GPU: a client running as a plugin inside a server
whose memory pool already owns the device, a loop that cannot afford the device synchronisation hidden
in cudaMalloc/cudaFree, or a plain C / ctypes driver that only hands over input and output buffers.
For them ``cs_tiny_ecg_train_step_noalloc`` keeps the gradient scratch in a built-in device array.
This test drives it with ctypes exactly as such a caller would and checks it against the regular
entry point (caller-provided scratch) on two consecutive steps, for the fp32 and the bf16 kernels.

The file name sorts last on purpose: a run that stops at the first failure has then already run
every other test.  It also passes on its own in a fresh process.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
STEP_ARGS = [vp, vp, vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, f32, f32, f32, i32, i32, f32, f32, i32, vp]
NOALLOC_ARGS = [vp, vp, vp, i32, i32, i32, i32, vp, vp, vp, vp, f32, f32, f32, i32, i32, f32, f32, i32, vp]


def _lib():
    from crossscale.utils import build
    import os
    path = build.cuda_lib_path()
    if not os.path.exists(path):
        build.build_cuda()
    lib = C.CDLL(path)
    lib.cs_last_error.restype = C.c_char_p
    lib.cs_tiny_ecg_train_step.restype = i32
    lib.cs_tiny_ecg_train_step.argtypes = STEP_ARGS
    lib.cs_tiny_ecg_train_step_noalloc.restype = i32
    lib.cs_tiny_ecg_train_step_noalloc.argtypes = NOALLOC_ARGS
    lib.cs_tiny_ecg_scratch_floats.restype = C.c_size_t
    lib.cs_tiny_ecg_scratch_floats.argtypes = [i32]
    return lib


def _state(params0, dev):
    return {"params": params0.clone(), "mom": torch.zeros_like(params0),
            "ctrl": torch.zeros(8, dtype=torch.int32, device=dev), "fstate": torch.zeros(8, device=dev)}


@pytest.mark.parametrize("amp", [0, 1])          # CS_AMP_FP32, CS_AMP_BF16
def test_noalloc_step_matches_regular_entry_point_twice(amp):
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(123)
    ncls, batch, length = 5, 96, 360
    nparam = 1424 + 17 * ncls
    params0 = torch.randn(nparam + 3, device=dev)[:nparam].contiguous() * 0.2
    params0 = torch.nn.functional.pad(params0, (0, (-nparam) % 4))
    data = torch.randn(4096, length, device=dev)
    labels = torch.randint(0, ncls, (4096,), device=dev)
    ref, new = _state(params0, dev), _state(params0, dev)
    ref["fstate"][1] = new["fstate"][1] = 1.0
    scratch = torch.zeros(int(lib.cs_tiny_ecg_scratch_floats(ncls)), device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    hyper = (0.05, 0.9, 0.0, amp, 0, 2.0, 0.5, 2000)
    for call in range(2):                                    # two calls in a row share the built-in array
        idx = torch.randint(0, 4096, (batch,), device=dev)
        rc = lib.cs_tiny_ecg_train_step(data.data_ptr(), labels.data_ptr(), idx.data_ptr(), 0, batch, length, ncls,
                                        ref["params"].data_ptr(), ref["mom"].data_ptr(), scratch.data_ptr(),
                                        ref["ctrl"].data_ptr(), ref["fstate"].data_ptr(), *hyper, stream)
        assert rc == 0, lib.cs_last_error()
        rc = lib.cs_tiny_ecg_train_step_noalloc(data.data_ptr(), labels.data_ptr(), idx.data_ptr(), 0, batch, length,
                                                ncls, new["params"].data_ptr(), new["mom"].data_ptr(),
                                                new["ctrl"].data_ptr(), new["fstate"].data_ptr(), *hyper, stream)
        assert rc == 0, lib.cs_last_error()
        torch.cuda.synchronize()
        # same kernel, same inputs: only the order of the floating-point gradient reductions differs
        torch.testing.assert_close(new["params"], ref["params"], rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(new["mom"], ref["mom"], rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(new["fstate"], ref["fstate"], rtol=1e-4, atol=1e-6)
        assert torch.equal(new["ctrl"], ref["ctrl"]) and int(new["ctrl"][2]) == call + 1
    assert not torch.equal(new["params"], params0)           # it did train


def test_noalloc_rejects_more_classes_than_the_array_holds():
    lib = _lib()
    dev = torch.device("cuda", 0)
    buf = torch.zeros(4096, device=dev)
    lab = torch.zeros(16, dtype=torch.int64, device=dev)
    ctrl = torch.zeros(8, dtype=torch.int32, device=dev)
    rc = lib.cs_tiny_ecg_train_step_noalloc(buf.data_ptr(), lab.data_ptr(), None, 0, 4, 64, 9, buf.data_ptr(),
                                            buf.data_ptr(), ctrl.data_ptr(), buf.data_ptr(), 0.01, 0.9, 0.0, 0, 0,
                                            2.0, 0.5, 2000, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.cs_last_error().decode()
    assert "9 classes" in msg and "7168" in msg


@pytest.mark.parametrize("amp", [0, 1])
def test_noalloc_step_clears_what_an_earlier_caller_left_in_the_array(amp):
    """The built-in array outlives every call.  Fill it with NaNs (what a torn-down caller, or a step with more
    classes, may leave) and check that the next step still matches the regular entry point: the entry point
    clears the part it uses before it launches."""
    lib = _lib()
    lib.cs_debug_noalloc_poison.restype = i32
    lib.cs_debug_noalloc_poison.argtypes = [vp]
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(7)
    ncls, batch, length = 2, 64, 500
    nparam = 1424 + 17 * ncls
    params0 = torch.nn.functional.pad(torch.randn(nparam, device=dev) * 0.2, (0, (-nparam) % 4))
    data = torch.randn(512, length, device=dev)
    labels = torch.randint(0, ncls, (512,), device=dev)
    ref, new = _state(params0, dev), _state(params0, dev)
    ref["fstate"][1] = new["fstate"][1] = 1.0
    scratch = torch.zeros(int(lib.cs_tiny_ecg_scratch_floats(ncls)), device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    hyper = (0.05, 0.9, 0.0, amp, 0, 2.0, 0.5, 2000)
    idx = torch.randint(0, 512, (batch,), device=dev)
    assert lib.cs_debug_noalloc_poison(stream) == 0, lib.cs_last_error()
    rc = lib.cs_tiny_ecg_train_step(data.data_ptr(), labels.data_ptr(), idx.data_ptr(), 0, batch, length, ncls,
                                    ref["params"].data_ptr(), ref["mom"].data_ptr(), scratch.data_ptr(),
                                    ref["ctrl"].data_ptr(), ref["fstate"].data_ptr(), *hyper, stream)
    assert rc == 0, lib.cs_last_error()
    rc = lib.cs_tiny_ecg_train_step_noalloc(data.data_ptr(), labels.data_ptr(), idx.data_ptr(), 0, batch, length, ncls,
                                            new["params"].data_ptr(), new["mom"].data_ptr(), new["ctrl"].data_ptr(),
                                            new["fstate"].data_ptr(), *hyper, stream)
    assert rc == 0, lib.cs_last_error()
    torch.cuda.synchronize()
    assert torch.isfinite(new["params"]).all() and int(new["ctrl"][3]) == 0      # no skipped step: nothing non-finite got in
    torch.testing.assert_close(new["params"], ref["params"], rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(new["mom"], ref["mom"], rtol=1e-4, atol=1e-6)
