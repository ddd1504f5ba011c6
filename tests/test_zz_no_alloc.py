"""pcv_attn_bwd_static — the allocation-free variant of pcv_attn_bwd (its scratch, delta and the fp32 dQ accumulator, is a
file-scope device array inside the library, no counterpart in the
reference) — called through the C ABI and held to the existing entry point on two calls in a row.

The file name sorts last on purpose: a failure here must not hide other tests.
Tolerance: dK and dV are accumulated in tensor memory in a fixed order, so they must be bit-identical; dQ is summed by
fp32 reduce-adds whose order differs from launch to launch, which moves an occasional bf16 rounding: rel. L2 <= 2e-3
(the kernel-level attention tests hold each of them to 1.5e-2 of the fp32 oracle)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _doc_starts(B, T, mean, seed):
    g = torch.Generator().manual_seed(seed)
    ds = torch.zeros(B, T, dtype=torch.int32)
    for b in range(B):
        t = 0
        while t < T:
            n = int(max(1, min(T - t, torch.empty(1).exponential_(1.0 / mean, generator=g).item())))
            ds[b, t:t + n] = t
            t += n
    return ds


def _case(B, T, NH, doc_mean, seed):
    from plaincv_b200 import ops
    g = torch.Generator().manual_seed(seed)
    d = NH * 64
    qkv = torch.randn(B * T, 3 * d, generator=g).to(torch.bfloat16).cuda()
    dout = torch.randn(B * T, d, generator=g).to(torch.bfloat16).cuda()
    cos_sin = ops.rope_table(64, T, 500000.0, "cuda")
    ds = None if doc_mean is None else _doc_starts(B, T, doc_mean, seed).cuda().view(-1)
    qkv = ops.rope_qk_(qkv, cos_sin, B, T, NH)
    out, lse = ops.attn_fwd(qkv, ds, B, T, NH)
    return qkv, cos_sin, ds, out, dout, lse


def _static_bwd(lib, qkv, cos_sin, ds, out, dout, lse, dqkv, B, T, NH):
    """The raw C-ABI call: pointers, sizes and the current stream; the output buffer exists already."""
    return lib.pcv_attn_bwd_static(qkv.data_ptr(), cos_sin.data_ptr(), None if ds is None else ds.data_ptr(),
                                   out.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(), B, T, NH,
                                   torch.cuda.current_stream().cuda_stream)


def test_attn_bwd_static_matches_attn_bwd_on_two_calls_in_a_row():
    from plaincv_b200 import _lib, ops
    lib = _lib.load()
    # call 1: causal, the largest problem the built-in scratch takes at 16 heads (B*T*d = 4 Mi floats);
    # call 2: a smaller, document-masked problem right behind it, so it finds call 1's contents in the scratch
    for k, (B, T, NH, doc_mean) in enumerate([(2, 2048, 16, None), (3, 384, 4, 90)]):
        qkv, cos_sin, ds, out, dout, lse = _case(B, T, NH, doc_mean, seed=40 + k)
        d = NH * 64
        ref = ops.attn_bwd(qkv, cos_sin, ds, out, dout, lse, B, T, NH)  # existing entry point, torch-allocated scratch
        got = torch.full_like(ref, float("nan"))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = _static_bwd(lib, qkv, cos_sin, ds, out, dout, lse, got, B, T, NH)
        torch.cuda.synchronize()
        assert rc == 0, rc
        assert torch.cuda.mem_get_info()[0] == free_before  # the call allocated no device memory
        assert torch.isfinite(got.float()).all()
        assert torch.equal(got[:, d:], ref[:, d:]), f"call {k}: dK / dV differ"
        e = rel_l2(got[:, :d], ref[:, :d])
        print(f"call {k}: B={B} T={T} heads={NH} doc_mean={doc_mean}: dQ rel-L2 vs pcv_attn_bwd {e:.2e}, dK/dV bit-identical")
        assert e <= 2e-3, (k, e)


def test_attn_bwd_static_rejects_what_does_not_fit():
    from plaincv_b200 import _lib
    lib = _lib.load()
    B, T, NH = 2, 256, 4
    qkv, cos_sin, ds, out, dout, lse = _case(B, T, NH, None, seed=50)
    dqkv = torch.zeros_like(qkv)
    sentinel = dqkv.clone()
    # 4 x 2048 tokens at 16 heads = 8 Mi floats of dQ accumulator: twice the limit -> PCV_ERR_SCRATCH, nothing launched
    n0 = lib.pcv_launch_count()
    rc = _static_bwd(lib, qkv, cos_sin, ds, out, dout, lse, dqkv, 4, 2048, 16)
    assert rc == -5
    assert lib.pcv_launch_count() == n0
    torch.cuda.synchronize()
    assert torch.equal(dqkv, sentinel)
    # argument errors are those of pcv_attn_bwd
    assert _static_bwd(lib, qkv, cos_sin, ds, out, dout, lse, dqkv, B, 200, NH) == -1  # T not a multiple of 128
