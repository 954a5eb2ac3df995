"""Which shared-memory word feeds which (row, k) of a tcgen05 operand?  One MMA on index-coded shared-memory images
(`cs_debug_mma_layout_probe`, csrc/cuda/mma_probe.cu) pins down the operand layouts the fused TinyECG kernels rely on,
measured on the device instead of read off the documentation.  No reference counterpart (the reference has no tensor-core
code).  Needs a B200."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lib():
    from crossscale.ops import _lib
    if not _lib.is_sm100():
        pytest.skip("needs sm_100")
    lib = _lib.cuda_lib()
    lib.cs_debug_mma_layout_probe.argtypes = [ctypes.c_int] * 9 + [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                                                  ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    return lib


def _run(fmt, M, N, amn, bmn, lboa, sboa, lbob, sbob, a_img, b_img, la=0, lb=0):
    out = np.zeros((128, N), dtype=np.float32)
    a_img, b_img = np.ascontiguousarray(a_img), np.ascontiguousarray(b_img)
    rc = _lib().cs_debug_mma_layout_probe(fmt, M, N, amn, bmn, lboa, sboa, lbob, sbob, a_img.ctypes.data, a_img.nbytes,
                                          b_img.ctypes.data, b_img.nbytes, out.ctypes.data, la, lb)
    assert rc == 0
    return out if M == 128 else np.concatenate([out[q * 32:q * 32 + 16] for q in range(4)])   # M = 64: 16 lanes per quadrant


def _ident_k_tf32():
    """K-major tf32 operand (LBO 256, SBO 128) with X[r, k] = (r == k), r, k < 8"""
    img = np.zeros(1024, dtype=np.float32)
    for n in range(8):
        img[((n >> 2) * 256 + (n & 7) * 16 + (n & 3) * 4) // 4] = 1.0
    return img


WORDS = np.arange(1, 2049, dtype=np.float32)        # word w holds w + 1: exact in tf32 up to 2048


def test_tf32_k_major_no_swizzle():
    """element (r, k) at (r/8)*SBO + (k/4)*LBO + (r%8)*16 + (k%4)*4: 16-byte rows hold FOUR 32-bit values"""
    d = _run(2, 64, 16, 0, 0, 4096, 128, 256, 128, WORDS, _ident_k_tf32())
    want = np.array([[((m // 8) * 128 + (k // 4) * 4096 + (m % 8) * 16 + (k % 4) * 4) // 4 + 1 for k in range(8)] for m in range(64)])
    assert np.array_equal(d[:, :8], want.astype(np.float32))


def test_tf32_mn_major_no_swizzle_reads_nothing():
    """the interleaved (no-swizzle) MN-major form is not available to 32-bit operands: the MMA returns zeros, no fault"""
    for lbo, sbo in [(128, 16), (16, 128)]:
        d = _run(2, 64, 16, 1, 0, lbo, sbo, 256, 128, np.ones(2048, np.float32), np.ones(1024, np.float32))
        assert not d.any()


def _swz(addr):
    return addr ^ ((addr >> 2) & 0x60)


@pytest.mark.parametrize("lbo,sbo", [(1024, 512), (512, 1024), (128, 512), (256, 512)])
def test_tf32_mn_major_swizzle128_base32(lbo, sbo):
    """layout type 1: logical address (mn/32)*LBO + (k/4)*SBO + (k%4)*128 + (mn%32)*4, bits 7..8 xor into bits 5..6 of the
    ABSOLUTE address - so LBO = 128 (the next atom is the array shifted by one row: the conv-tap trick) just works."""
    d = _run(2, 64, 16, 1, 0, lbo, sbo, 256, 128, WORDS, _ident_k_tf32(), la=1)
    want = np.array([[_swz((m // 32) * lbo + (k // 4) * sbo + (k % 4) * 128 + (m % 32) * 4) // 4 + 1 for k in range(8)]
                     for m in range(64)], dtype=np.float32)
    assert np.array_equal(d[:, :8], want)


def test_tf32_mn_major_b_operand_32_wide():
    a = np.zeros(2048, np.float32)
    a[:1024] = _ident_k_tf32()
    d = _run(2, 64, 32, 0, 1, 256, 128, 1024, 512, a, WORDS, lb=1)
    want = np.array([[_swz((k // 4) * 512 + (k % 4) * 128 + n * 4) // 4 + 1 for k in range(8)] for n in range(32)], dtype=np.float32)
    assert np.array_equal(d[:8, :32].T, want)


def test_bf16_mn_major_no_swizzle_shifted_groups():
    """16-bit MN-major, no swizzle: element (mn, k) at (mn/8)*SBO + (k/8)*LBO + (k%8)*16 + (mn%8)*2; with SBO = 16 the
    M-group g is the plane shifted by g rows (conv taps in the bf16 / fp16 weight-gradient MMAs)"""
    import torch
    vals = (torch.arange(2048) % 256 + 1).to(torch.float32)            # value of 16-bit element e: e % 256 + 1 (exact in bf16)
    a = vals.to(torch.bfloat16).view(torch.int16).numpy()
    one = int(torch.tensor([1.0], dtype=torch.bfloat16).view(torch.int16).item())
    b = np.zeros(2048, dtype=np.int16)
    for n in range(16):
        b[((n >> 3) * 128 + (n >> 3) * 256 + (n & 7) * 16 + (n & 7) * 2) // 2] = one      # K-major B[n, k] = (n == k)
    d = _run(1, 64, 16, 1, 0, 128, 16, 256, 128, a.view(np.float32), b.view(np.float32))
    want = np.array([[(((m // 8) * 16 + (k // 8) * 128 + (k % 8) * 16 + (m % 8) * 2) // 2) % 256 + 1 for k in range(16)]
                     for m in range(64)], dtype=np.float32)
    assert np.array_equal(d, want)
