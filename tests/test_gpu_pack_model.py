"""GPU parity: K4/K5 ragged pack / unpack (bit-exact layout) and K6/K8 mock-model forwards."""
import numpy as np
import pytest
import torch

from oracle import models

pytestmark = pytest.mark.gpu

NP_OF = {1: np.uint8, 2: np.int16, 4: np.float32, 8: np.int64}


def ragged(rng, n_req, max_len, elem_bytes, min_len=0):
    lens = rng.integers(min_len, max_len + 1, n_req)
    off = np.zeros(n_req + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    total = int(off[-1])
    if elem_bytes == 4:
        vals = rng.standard_normal(total + 8).astype(np.float32)
    else:
        vals = rng.integers(0, 120, total + 8).astype(NP_OF[elem_bytes])
    return vals, off, lens


def pack_ref(vals, off, Lpad, src_index=None):
    idx = range(len(off) - 1) if src_index is None else src_index
    out = np.zeros((len(idx), Lpad), dtype=vals.dtype)
    for r, q in enumerate(idx):
        n = min(int(off[q + 1] - off[q]), Lpad)
        out[r, :n] = vals[off[q]:off[q] + n]
    return out


@pytest.mark.parametrize("elem_bytes,n_req,Lpad,min_len", [(4, 256, 2048, 16), (4, 1000, 64, 0), (4, 4096, 128, 128),
                                                           (1, 300, 48, 0), (2, 300, 40, 0), (8, 200, 10, 0),
                                                           (4, 77, 13, 0), (1, 50, 7, 0),
                                                           # rows of 1 KB and more: the bulk-copy kernels (csrc/pack.cu)
                                                           (4, 500, 256, 0), (8, 300, 512, 3), (4, 700, 4096, 1),
                                                           (1, 300, 2048, 0), (2, 300, 1024, 5), (4, 2000, 1024, 1000)])
def test_pack_layout_and_unpack_roundtrip(elem_bytes, n_req, Lpad, min_len, monkeypatch):
    from die import ops
    if n_req % 2 == 0:   # half of the shapes through the opt-in bulk-copy kernels (rows of 1 KB and more take them)
        monkeypatch.setenv("DIE_PACK_BULK", "1")
    rng = np.random.default_rng(n_req + Lpad)
    vals, off, lens = ragged(rng, n_req, Lpad, elem_bytes, min_len)
    v_d, off_d = torch.from_numpy(vals).cuda(), torch.from_numpy(off).cuda()
    packed, lengths = ops.pack_ragged(v_d, off_d, Lpad)
    assert np.array_equal(packed.cpu().numpy(), pack_ref(vals, off, Lpad))      # row i = request i, zero padded
    assert lengths.cpu().tolist() == lens.tolist()
    # gather form: row r = request src_index[r] (the misses of a batch, in arrival order)
    sel = np.sort(rng.choice(n_req, size=max(n_req // 3, 1), replace=False)).astype(np.int32)
    packed_sel, lengths_sel = ops.pack_ragged(v_d, off_d, Lpad, torch.from_numpy(sel).cuda())
    assert np.array_equal(packed_sel.cpu().numpy(), pack_ref(vals, off, Lpad, sel))
    assert lengths_sel.cpu().tolist() == lens[sel].tolist()
    # unpack is the exact inverse on the payload bytes; bytes outside the slices stay untouched
    sentinel = 77 if elem_bytes != 4 else -5.0
    flat = torch.full((len(vals),), sentinel, dtype=v_d.dtype, device="cuda")
    ops.unpack_ragged(packed, off_d, flat)
    got = flat.cpu().numpy()
    assert np.array_equal(got[:off[-1]], vals[:off[-1]]) and np.all(got[off[-1]:] == sentinel)
    # scatter form: packed row r -> request dst_index[r]
    flat2 = torch.full((len(vals),), sentinel, dtype=v_d.dtype, device="cuda")
    ops.unpack_ragged(packed_sel, off_d, flat2, torch.from_numpy(sel).cuda())
    want = np.full(len(vals), sentinel, dtype=vals.dtype)
    for q in sel:
        want[off[q]:off[q + 1]] = vals[off[q]:off[q + 1]]
    assert np.array_equal(flat2.cpu().numpy(), want)


def test_pack_truncates_rows_longer_than_lpad_and_handles_empty():
    from die import ops
    vals = np.arange(100, dtype=np.float32)
    off = np.array([0, 0, 30, 30, 100], dtype=np.int64)
    packed, lengths = ops.pack_ragged(torch.from_numpy(vals).cuda(), torch.from_numpy(off).cuda(), 32)
    assert lengths.cpu().tolist() == [0, 30, 0, 70]
    assert np.array_equal(packed.cpu().numpy(), pack_ref(vals, off, 32))
    empty, _ = ops.pack_ragged(torch.zeros(4, device="cuda"), torch.zeros(1, dtype=torch.int64, device="cuda"), 8)
    assert empty.shape == (0, 8)


def test_pack_unpack_full_size_roundtrip_property():
    """configs[2] shape: 64 flushes x 256 requests of 16..2048 fp32 -> pack -> unpack == identity."""
    from die import ops
    rng = np.random.default_rng(2)
    vals, off, lens = ragged(rng, 256 * 64, 2048, 4, 16)
    v_d, off_d = torch.from_numpy(vals).cuda(), torch.from_numpy(off).cuda()
    packed, lengths = ops.pack_ragged(v_d, off_d, 2048)
    flat = torch.zeros_like(v_d)
    ops.unpack_ragged(packed, off_d, flat)
    assert torch.equal(flat[:off[-1]], v_d[:off[-1]])
    assert int(lengths.sum()) == int(off[-1])
    # checksum of the padded batch == checksum of the ragged payload (padding is all zero)
    assert int(packed.view(torch.int32).to(torch.int64).sum()) == int(v_d[:off[-1]].view(torch.int32).to(torch.int64).sum())


def test_echo_is_bit_exact():
    from die import ops
    for n in (1, 5, 4096 * 128, 1000003):
        x = torch.randn(n, device="cuda")
        assert torch.equal(ops.echo_fwd(x), x)
    xb = torch.randint(0, 255, (1001,), dtype=torch.uint8, device="cuda")
    assert torch.equal(ops.echo_fwd(xb[1:]), xb[1:])  # unaligned bytes


@pytest.mark.parametrize("B,K,N", [(4096, 128, 128), (1, 128, 128), (63, 100, 72), (1000, 2048, 256)])
def test_linear_fp32_within_tolerance_of_f64_oracle(B, K, N):
    """Tolerance (oracle/models.py): |d| <= 1e-5 * sum_k |x_k w_k| + 1e-6 for the fp32 FMA path."""
    from die import ops
    rng = np.random.default_rng(B + K)
    x = rng.standard_normal((B, K)).astype(np.float32)
    w = (rng.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
    b = rng.standard_normal(N).astype(np.float32)
    y = ops.linear_fwd(torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda()).cpu().numpy()
    want = models.linear_f64(x, w, b)
    tol = 1e-5 * models.linear_abs_bound(x, w) + 1e-6
    assert np.all(np.abs(y.astype(np.float64) - want) <= tol)


@pytest.mark.parametrize("B,K,N", [(4096, 128, 128), (1, 128, 128), (129, 32, 16), (1000, 64, 256), (3670, 128, 128), (300, 256, 64),
                                   (37950, 128, 128), (40001, 64, 64), (38000, 128, 32)])  # the last three: streaming form
def test_linear_tf32_tensor_core_within_tolerance(B, K, N):
    """tcgen05 kind::tf32 path.  Tolerance (oracle/models.py): |d| <= 2**-10 * sum_k |x_k w_k| + 1e-6."""
    from die import _cabi, ops
    rng = np.random.default_rng(B * 3 + K)
    x = rng.standard_normal((B, K)).astype(np.float32)
    w = (rng.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
    b = rng.standard_normal(N).astype(np.float32)
    xd, wd, bd = torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda()
    y = ops.linear_fwd(xd, wd, bd, _cabi.LINEAR_TF32).cpu().numpy()
    want = models.linear_f64(x, w, b)
    tol = 2.0 ** -10 * models.linear_abs_bound(x, w) + 1e-6
    err = np.abs(y.astype(np.float64) - want)
    assert np.all(err <= tol), (float(err.max()), float(tol.min()))
    # and it is not the fp32 kernel in disguise: tf32 truncation must be visible somewhere
    y32 = ops.linear_fwd(xd, wd, bd, _cabi.LINEAR_FP32).cpu().numpy()
    assert B * N < 2000 or not np.array_equal(y, y32)


def test_linear_tf32_rejects_unsupported_shapes():
    from die import _cabi, ops
    x = torch.zeros((8, 100), device="cuda")
    w = torch.zeros((72, 100), device="cuda")
    with pytest.raises(_cabi.DieError):
        ops.linear_fwd(x, w, None, _cabi.LINEAR_TF32)


@pytest.mark.parametrize("B,D,H", [(256, 512, 512), (100, 256, 1024), (1024, 4096, 4096)])
def test_mlp_bf16_tensor_core_within_tolerance(B, D, H):
    """configs[4] MLP on tcgen05 (TMA + TMEM).  Tolerance (oracle/models.py): per row
    |d| <= 2**-7 * |y| + 1e-2 * max_j |y_j|  (bf16 output rounding + fp32 accumulation order)."""
    from die import ops
    rng = np.random.default_rng(B + D)
    x = models.to_bf16(rng.standard_normal((B, D)).astype(np.float32))
    w1 = models.to_bf16((rng.standard_normal((H, D)) / np.sqrt(D)).astype(np.float32))
    w2 = models.to_bf16((rng.standard_normal((D, H)) / np.sqrt(H)).astype(np.float32))
    b1 = rng.standard_normal(H).astype(np.float32)
    b2 = rng.standard_normal(D).astype(np.float32)
    dev = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    y = ops.mlp_fwd_bf16(dev(x).bfloat16(), dev(w1).bfloat16(), dev(b1), dev(w2).bfloat16(), dev(b2))
    y = y.float().cpu().numpy().astype(np.float64)
    want = models.mlp_bf16(x, w1, b1, w2, b2)
    tol = 2.0 ** -7 * np.abs(want) + 1e-2 * np.abs(want).max(axis=1, keepdims=True)
    err = np.abs(y - want)
    assert np.all(err <= tol), (float(err.max()), float((err / tol).max()))
    assert float(np.abs(y).mean()) > 0.1  # not all zeros
