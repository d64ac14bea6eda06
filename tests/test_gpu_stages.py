"""Stage-level parity on a real B200: every kernel is called through the C-ABI
(ctypes -> libspu_b200.so) and compared with the CPU oracle / an fp32 restatement of
the same op on the same seeded inputs.  Integer/byte stages must be bit-exact."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def spu(built_libs):
    from starpu_inference_server_b200 import capi

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    L = capi.lib()
    assert L.spu_device_is_sm100(0) == 1, "device 0 is not sm_100"
    return capi


def _sync():
    torch.cuda.synchronize()


# --------------------------------------------------------------------------- GEMM
def _gemm_ref(a_bf16, w_bf16, bias, resid, epi):
    y = a_bf16.float().cpu() @ w_bf16.float().cpu().T + bias.cpu()
    if epi == 1:
        y = torch.nn.functional.gelu(y)  # exact erf GELU
    if epi == 2:
        y = y + resid.cpu()
    return y


def _check_gemm(spu, M, N, K, epi, cg, bn, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    a = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).cuda()
    w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).cuda()
    bias = torch.randn(N, generator=g).cuda()
    resid = torch.randn(M, N, generator=g).cuda()
    # a guard row after the output catches stores that ignore the M tail
    out = torch.full((M + 1, N), float("nan"), dtype=torch.float32 if epi == 2 else torch.bfloat16, device="cuda")
    spu.check(spu.lib().spu_gemm_bf16_cfg_async(spu.ptr(a), spu.ptr(w), spu.ptr(bias),
                                                spu.ptr(resid) if epi == 2 else None, spu.ptr(out),
                                                M, N, K, epi, cg, bn, spu.stream_ptr()))
    _sync()
    ref = _gemm_ref(a, w, bias, resid, epi)
    assert torch.isnan(out[M]).all(), "GEMM wrote past row M"
    got = out[:M].float().cpu()
    assert torch.isfinite(got).all()
    # fp32 accumulation of exact bf16 products: only summation order + output rounding differ
    tol = 2e-3 if epi == 2 else 1.2e-2  # bf16 output rounding (2^-8 relative) dominates for epi 0/1
    err = (got - ref).abs()
    bound = tol * (ref.abs() + 1.0)
    bad = (err > bound).sum().item()
    assert bad == 0, f"{bad} elements off; max err {err.max().item():.4g} at {np.unravel_index(err.argmax().item(), err.shape)}"


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (128, 256, 128), (256, 768, 768), (4096, 2304, 768),
                                   (4096, 768, 3072), (1000, 3072, 768), (77, 384, 128), (640, 512, 128)])
@pytest.mark.parametrize("epi", [0, 1, 2])
def test_gemm_tcgen05_matches_fp32(spu, M, N, K, epi):
    """Library-chosen tile (cta_group, BLOCK_N)."""
    _check_gemm(spu, M, N, K, epi, 0, 0, M * 7 + N * 3 + K + epi)


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("bn", [64, 128, 192, 256])
@pytest.mark.parametrize("epi", [0, 1, 2])
@pytest.mark.parametrize("M", [4096, 333, 31])
def test_gemm_every_tile_shape(spu, M, epi, bn, cg):
    """Every (cta_group, BLOCK_N) instantiation, with full, ragged and sub-slab M; N=768
    gives 3..12 column blocks and K=320 five k-blocks (pipeline wrap with <= 8 stages
    needs several tiles per CTA: M=4096 supplies them)."""
    _check_gemm(spu, M, 768, 320 if M != 4096 else 768, epi, cg, bn, M + 13 * bn + cg + epi)


@pytest.mark.parametrize("bn", [128, 192, 256])
@pytest.mark.parametrize("epi", [0, 1, 2])
@pytest.mark.parametrize("M,K", [(4096, 768), (4096, 3072), (333, 320), (31, 320), (8192, 768)])
def test_gemm_half_footprint_tiles(spu, M, K, epi, bn):
    """cta_group 3 = CTA pair of half-footprint CTAs (two resident per SM: 3-4 ring stages, 4 epilogue warps, one
    accumulator buffer above 128 columns).  K=3072 wraps the short ring 16 times, M=8192 gives CTAs several tiles."""
    if bn == 256 and epi == 2:  # 3 x 32 KB stages + fp32 staging exceed 113 KB
        with pytest.raises(Exception, match="footprint"):
            _check_gemm(spu, M, 768, K, epi, 3, bn, 1)
        return
    _check_gemm(spu, M, 768, K, epi, 3, bn, M + 13 * bn + 3 + epi + K)


def test_gemm_half_footprint_inplace_residual(spu):
    M, N, K = 4096, 768, 3072
    g = torch.Generator().manual_seed(6)
    a = torch.randn(M, K, generator=g).to(torch.bfloat16).cuda()
    w = (torch.randn(N, K, generator=g) * 0.02).to(torch.bfloat16).cuda()
    bias = torch.randn(N, generator=g).cuda()
    x = torch.randn(M, N, generator=g).cuda()
    ref = _gemm_ref(a, w, bias, x, 2)
    spu.check(spu.lib().spu_gemm_bf16_cfg_async(spu.ptr(a), spu.ptr(w), spu.ptr(bias), spu.ptr(x), spu.ptr(x),
                                                M, N, K, 2, 3, 192, spu.stream_ptr()))
    _sync()
    assert torch.allclose(x.cpu(), ref, rtol=2e-3, atol=2e-3)


def test_gelu_epilogue_tracks_erf(spu):
    """The 2^p(|x|) GELU of the epilogue against exact erf over a wide input range
    (identity weights route A straight into the activation)."""
    K = N = 256
    M = 4096
    x = torch.linspace(-9.0, 9.0, M * K).reshape(M, K)
    a = x.to(torch.bfloat16).cuda()
    w = torch.eye(N, K).to(torch.bfloat16).cuda()
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    spu.check(spu.lib().spu_gemm_bf16_async(spu.ptr(a), spu.ptr(w), None, None, spu.ptr(out), M, N, K, 1,
                                            spu.stream_ptr()))
    _sync()
    xin = a.float().cpu().double()
    ref = 0.5 * xin * (1.0 + torch.erf(xin / 2 ** 0.5))
    err = (out.float().cpu().double() - ref).abs()
    assert (err <= 2.0 ** -8 * ref.abs() + 1e-5).all(), err.max().item()  # one bf16 rounding + 1e-5


def test_gemm_inplace_residual(spu):
    M, N, K = 512, 768, 768
    g = torch.Generator().manual_seed(5)
    a = torch.randn(M, K, generator=g).to(torch.bfloat16).cuda()
    w = (torch.randn(N, K, generator=g) * 0.03).to(torch.bfloat16).cuda()
    bias = torch.randn(N, generator=g).cuda()
    x = torch.randn(M, N, generator=g).cuda()
    ref = _gemm_ref(a, w, bias, x, 2)
    spu.check(spu.lib().spu_gemm_bf16_async(spu.ptr(a), spu.ptr(w), spu.ptr(bias), spu.ptr(x), spu.ptr(x),
                                            M, N, K, 2, spu.stream_ptr()))
    _sync()
    assert torch.allclose(x.cpu(), ref, rtol=2e-3, atol=2e-3)


# ------------------------------------------- GEMM with fused residual + LayerNorm epilogue
def _check_gemm_ln(spu, M, N, K, cg, bn, seed, inplace=True, want_f32=True, want_bf16=True):
    g = torch.Generator(device="cpu").manual_seed(seed)
    a = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).cuda()
    w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).cuda()
    bias, gamma, beta = (torch.randn(N, generator=g).cuda() for _ in range(3))
    # a non-zero row mean and unequal column scales: the merged (mean, M2) must handle both
    resid = (torch.randn(M, N, generator=g) * torch.linspace(0.2, 3.0, N) + 1.5).cuda()
    x = a.float().cpu() @ w.float().cpu().T + bias.cpu() + resid.cpu()
    ref = torch.nn.functional.layer_norm(x.double(), (N,), gamma.cpu().double(), beta.cpu().double(), 1e-12).float()
    guard = torch.full((M + 1, N), float("nan"), device="cuda")
    if inplace:
        guard[:M] = resid
        resid_arg = guard
    else:
        resid_arg = resid
    out16 = torch.full((M + 1, N), float("nan"), dtype=torch.bfloat16, device="cuda")
    spu.check(spu.lib().spu_gemm_ln_bf16_async(
        spu.ptr(a), spu.ptr(w), spu.ptr(bias), spu.ptr(resid_arg), spu.ptr(gamma), spu.ptr(beta), 1e-12,
        spu.ptr(guard) if want_f32 else None, spu.ptr(out16) if want_bf16 else None, M, N, K, cg, bn,
        spu.stream_ptr()))
    _sync()
    if want_f32:
        assert torch.isnan(guard[M]).all(), "fused LayerNorm GEMM wrote fp32 past row M"
        got = guard[:M].cpu()
        assert torch.isfinite(got).all()
        err = (got - ref).abs().max().item()
        assert err < 5e-4, f"fp32 y: max err {err:.3g}"
    if want_bf16:
        assert torch.isnan(out16[M].float()).all(), "fused LayerNorm GEMM wrote bf16 past row M"
        got16 = out16[:M].cpu()
        if want_f32:  # the bf16 copy is the fp32 result rounded to nearest even, bit for bit
            assert torch.equal(got16.view(torch.int16), guard[:M].cpu().to(torch.bfloat16).view(torch.int16))
        else:
            assert ((got16.float() - ref).abs() <= 2.0 ** -8 * ref.abs() + 5e-4).all()


@pytest.mark.parametrize("M,N,K,cg,bn", [
    (4096, 768, 768, 0, 0),     # BERT-base attention-output: 16 clusters of 4 CTA pairs x 192 columns
    (4096, 768, 3072, 0, 0),    # BERT-base FFN2
    (5000, 768, 768, 0, 0),     # more row blocks than resident clusters: persistent loop, ragged tail
    (333, 768, 320, 0, 0),
    (64, 768, 768, 0, 0),       # a single slab: the second CTA of each pair holds dead rows
    (1024, 768, 768, 2, 192),
    (512, 128, 128, 1, 64), (512, 128, 128, 2, 64), (512, 128, 128, 1, 128), (300, 128, 512, 2, 128),
    (700, 256, 128, 2, 64), (700, 256, 128, 1, 64), (700, 256, 256, 2, 128),
    (1111, 512, 128, 2, 128), (1111, 512, 192, 1, 128), (8192, 384, 128, 2, 192),
])
def test_gemm_fused_residual_layernorm(spu, M, N, K, cg, bn):
    """y = LayerNorm(A W^T + bias + resid) with the row statistics merged across the cluster that spans the N
    columns (st.async over distributed shared memory), against a float64 LayerNorm of the fp32 product."""
    _check_gemm_ln(spu, M, N, K, cg, bn, seed=M + N + K + cg + bn)


def test_gemm_fused_layernorm_output_variants(spu):
    _check_gemm_ln(spu, 600, 768, 768, 0, 0, seed=1, inplace=False)
    _check_gemm_ln(spu, 600, 768, 768, 0, 0, seed=2, want_bf16=False)
    _check_gemm_ln(spu, 600, 768, 768, 0, 0, seed=3, want_f32=False, inplace=False)
    # hidden sizes no cluster of <= 8 CTAs spans are refused (the forward then keeps the separate LayerNorm kernel)
    a = torch.zeros(128, 1024, dtype=torch.bfloat16, device="cuda")
    st = spu.lib().spu_gemm_ln_bf16_async(spu.ptr(a), spu.ptr(a), spu.ptr(a), spu.ptr(a), spu.ptr(a), spu.ptr(a), 1e-12,
                                          spu.ptr(a), None, 128, 1024, 1024, 0, 0, spu.stream_ptr())
    assert st == spu.SPU_ERR_UNSUPPORTED
    # 768 = 6 x 128: more column blocks than the cluster holds
    st = spu.lib().spu_gemm_ln_bf16_async(spu.ptr(a), spu.ptr(a), spu.ptr(a), spu.ptr(a), spu.ptr(a), spu.ptr(a), 1e-12,
                                          spu.ptr(a), None, 128, 768, 768, 2, 128, spu.stream_ptr())
    assert st == spu.SPU_ERR_UNSUPPORTED


def test_gemm_rejects_bad_shapes(spu):
    a = torch.zeros(128, 100, dtype=torch.bfloat16, device="cuda")
    st = spu.lib().spu_gemm_bf16_async(spu.ptr(a), spu.ptr(a), None, None, spu.ptr(a), 128, 128, 100, 0,
                                       spu.stream_ptr())
    assert st == spu.SPU_ERR_UNSUPPORTED
    assert b"K must be" in spu.lib().spu_last_error()


# ---------------------------------------------------------------------- attention
def _attn_ref(qkv, mask_bias, B, S, heads):
    H = heads * 64
    x = qkv.float().cpu().view(B, S, 3, heads, 64)
    q, k, v = (x[:, :, i].transpose(1, 2) for i in range(3))
    s = q @ k.transpose(-1, -2) / 8.0 + mask_bias.cpu().view(B, 1, 1, S)
    return (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B * S, H)


_ATTN_IMPLS = [(1, "tcgen05"), (2, "mma.sync")]
_ATTN_SHAPES = [(2, 128, 12), (1, 64, 2), (3, 200, 4), (2, 512, 12), (5, 16, 2), (1, 37, 1), (2, 129, 3), (3, 197, 16),
                (2, 256, 2), (1, 389, 2), (1, 1, 1), (33, 128, 12)]


def _attn_inputs(B, S, heads):
    g = torch.Generator().manual_seed(B * 100 + S)
    H = heads * 64
    qkv = torch.randn(B * S, 3 * H, generator=g).to(torch.bfloat16).cuda()
    lens = torch.randint(1, S + 1, (B,), generator=g)
    keep = torch.arange(S)[None, :] < lens[:, None]
    keep[0] = True
    mask_bias = torch.where(keep, 0.0, -1e30).float().cuda()
    return qkv, mask_bias


@pytest.mark.parametrize("impl,_name", _ATTN_IMPLS)
@pytest.mark.parametrize("B,S,heads", _ATTN_SHAPES)
def test_attention_matches_fp32(spu, B, S, heads, impl, _name):
    H = heads * 64
    qkv, mask_bias = _attn_inputs(B, S, heads)
    ctx = torch.full((B * S, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    spu.check(spu.lib().spu_attention_cfg_async(spu.ptr(qkv), spu.ptr(mask_bias), spu.ptr(ctx), B, S, heads, impl, 0,
                                                1.0, spu.stream_ptr()))
    _sync()
    ref = _attn_ref(qkv, mask_bias, B, S, heads)
    got = ctx.float().cpu()
    assert torch.isfinite(got).all()
    # P is rounded to bf16 before PV and the output is bf16
    assert (got - ref).abs().max().item() < 3e-2, (got - ref).abs().max().item()


@pytest.mark.parametrize("B,S,heads", [(2, 128, 12), (3, 197, 16), (2, 512, 4)])
def test_attention_no_mask_pointer(spu, B, S, heads):
    # ViT path: mask_bias == NULL; rows of a neighbouring sequence must never leak into a ragged tile
    H = heads * 64
    qkv, _ = _attn_inputs(B, S, heads)
    ctx = torch.full((B * S, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    spu.check(spu.lib().spu_attention_cfg_async(spu.ptr(qkv), None, spu.ptr(ctx), B, S, heads, 1, 0, 1.0,
                                                spu.stream_ptr()))
    _sync()
    ref = _attn_ref(qkv, torch.zeros(B, S), B, S, heads)
    assert (ctx.float().cpu() - ref).abs().max().item() < 3e-2


@pytest.mark.parametrize("B,S,heads", [(2, 128, 12), (2, 300, 4)])
def test_attention_tcgen05_tracks_mma_sync(spu, B, S, heads):
    # two independent implementations of the same bf16 pipeline: agree to bf16 rounding of the output
    H = heads * 64
    qkv, mask_bias = _attn_inputs(B, S, heads)
    outs = []
    for impl in (1, 2):
        ctx = torch.zeros(B * S, H, dtype=torch.bfloat16, device="cuda")
        spu.check(spu.lib().spu_attention_cfg_async(spu.ptr(qkv), spu.ptr(mask_bias), spu.ptr(ctx), B, S, heads, impl,
                                                    0, 1.0, spu.stream_ptr()))
        _sync()
        outs.append(ctx.float().cpu())
    assert (outs[0] - outs[1]).abs().max().item() < 1.6e-2


@pytest.mark.parametrize("impl,_name", _ATTN_IMPLS)
@pytest.mark.parametrize("B,S,heads", [(2, 128, 12), (2, 200, 4)])
def test_attention_e4m3_output(spu, B, S, heads, impl, _name):
    H = heads * 64
    qkv, mask_bias = _attn_inputs(B, S, heads)
    ref = _attn_ref(qkv, mask_bias, B, S, heads)
    scale = ref.abs().max().item() / 448.0
    ctx = torch.zeros(B * S, H, dtype=torch.uint8, device="cuda")
    spu.check(spu.lib().spu_attention_cfg_async(spu.ptr(qkv), spu.ptr(mask_bias), spu.ptr(ctx), B, S, heads, impl, 1,
                                                1.0 / scale, spu.stream_ptr()))
    _sync()
    got = ctx.view(torch.float8_e4m3fn).float().cpu() * scale
    # e4m3 has 3 mantissa bits: half-ulp relative error 2^-4, plus the bf16 pipeline error
    err = (got - ref).abs()
    assert (err <= 0.0625 * ref.abs() + 3e-2).all(), err.max().item()


@pytest.mark.parametrize("impl,_name", _ATTN_IMPLS)
def test_attention_fully_masked_row_is_uniform(spu, impl, _name):
    # HF adds finfo.min to every key of a fully padded sequence -> softmax is uniform
    B, S, heads = 1, 128, 1
    g = torch.Generator().manual_seed(3)
    qkv = torch.randn(B * S, 192, generator=g).to(torch.bfloat16).cuda()
    mask_bias = torch.full((B, S), -1e30, device="cuda")
    ctx = torch.zeros(B * S, 64, dtype=torch.bfloat16, device="cuda")
    spu.check(spu.lib().spu_attention_cfg_async(spu.ptr(qkv), spu.ptr(mask_bias), spu.ptr(ctx), B, S, heads, impl, 0,
                                                1.0, spu.stream_ptr()))
    _sync()
    v_mean = qkv.float().cpu()[:, 128:].mean(0)
    assert (ctx.float().cpu() - v_mean[None]).abs().max().item() < 2e-2


@pytest.mark.parametrize("B,S,heads", [(256, 128, 12), (64, 512, 12), (64, 197, 16)])
def test_attention_full_size_properties(spu, B, S, heads):
    """BASELINE.json's full sizes, checked through size-independent properties instead of an O(B*S*S) oracle:
    (1) softmax rows sum to one: with every V row of a head equal to a constant vector c the output is c;
    (2) padded keys are inert: garbage in the K/V rows of hard-masked keys does not change any output row;
    (3) sequences are independent: permuting the batch permutes the output."""
    H = heads * 64
    g = torch.Generator().manual_seed(S + B)
    qkv = torch.randn(B, S, 3 * H, generator=g).to(torch.bfloat16)
    c = torch.randn(H, generator=g).to(torch.bfloat16)
    lens = torch.randint(1, S + 1, (B,), generator=g)
    lens[0] = S
    keep = torch.arange(S)[None, :] < lens[:, None]
    mask_bias = torch.where(keep, 0.0, -1e30).float().cuda()

    def run(x, mb):
        x = x.reshape(B * S, 3 * H).contiguous().cuda()
        ctx = torch.full((B * S, H), float("nan"), dtype=torch.bfloat16, device="cuda")
        spu.check(spu.lib().spu_attention_cfg_async(spu.ptr(x), spu.ptr(mb), spu.ptr(ctx), B, S, heads, 1, 0, 1.0,
                                                    spu.stream_ptr()))
        _sync()
        return ctx.cpu().view(B, S, H)

    const_v = qkv.clone()
    const_v[:, :, 2 * H:] = c
    out = run(const_v, mask_bias).float()
    assert torch.isfinite(out).all()
    # P is rounded to bf16 (rel 2^-9) while the row sum is kept in fp32
    assert (out - c.float()).abs().max().item() <= 2e-2 * max(1.0, c.float().abs().max().item())

    base = run(qkv, mask_bias)
    noisy = qkv.clone()
    pad = ~keep
    noisy[:, :, H:][pad] = (torch.randn(int(pad.sum()), 2 * H, generator=g) * 50).to(torch.bfloat16)
    assert torch.equal(run(noisy, mask_bias), base)

    perm = torch.randperm(B, generator=g)
    assert torch.equal(run(qkv[perm], mask_bias[perm.cuda()].contiguous()), base[perm])


# --------------------------------------------------------------- LayerNorm / embed
@pytest.mark.parametrize("rows,hidden", [(4096, 768), (33, 128), (1000, 1024), (1, 768)])
def test_layernorm_matches_oracle(spu, rows, hidden):
    from oracle.bert_oracle import _ln

    g = torch.Generator().manual_seed(rows + hidden)
    x = (torch.randn(rows, hidden, generator=g) * 3 + 0.5)
    gamma, beta = torch.randn(hidden, generator=g), torch.randn(hidden, generator=g)
    ref = _ln(x, gamma, beta, 1e-12)
    xd, gd, bd = x.cuda(), gamma.cuda(), beta.cuda()
    o32 = torch.empty_like(xd)
    o16 = torch.empty(rows, hidden, dtype=torch.bfloat16, device="cuda")
    spu.check(spu.lib().spu_layernorm_async(spu.ptr(xd), spu.ptr(gd), spu.ptr(bd), 1e-12, spu.ptr(o32),
                                            spu.ptr(o16), rows, hidden, spu.stream_ptr()))
    _sync()
    assert torch.allclose(o32.cpu(), ref, rtol=1e-4, atol=1e-4)
    assert torch.equal(o16.cpu(), o32.cpu().to(torch.bfloat16))
    # in-place variant
    spu.check(spu.lib().spu_layernorm_async(spu.ptr(xd), spu.ptr(gd), spu.ptr(bd), 1e-12, spu.ptr(xd),
                                            None, rows, hidden, spu.stream_ptr()))
    _sync()
    assert torch.equal(xd.cpu(), o32.cpu())


def test_embed_ln_and_mask_bias(spu):
    from oracle import bert_oracle as bo

    spec = bo.BERT_TINY
    w = bo.make_weights(spec, 0)
    ids, mask = bo.make_inputs(spec, 4, 48, seed=5, min_len=3)
    mask[1] *= 7  # non-binary mask words mean "keep" (reference warm-up feeds randint(0,10))
    _, layers = bo.bert_forward(w, spec, ids, mask, return_layers=True)
    t = {k: torch.from_numpy(v).cuda() for k, v in w.items() if k.startswith("embeddings")}
    B, S, H = 4, 48, spec.hidden
    x32 = torch.empty(B * S, H, device="cuda")
    x16 = torch.empty(B * S, H, dtype=torch.bfloat16, device="cuda")
    mb = torch.empty(B * S, device="cuda")
    ids_d, mask_d = torch.from_numpy(ids).cuda(), torch.from_numpy(mask).cuda()  # keep alive
    spu.check(spu.lib().spu_bert_embed_async(
        spu.ptr(ids_d), spu.ptr(mask_d),
        spu.ptr(t["embeddings.word_embeddings.weight"]), spu.ptr(t["embeddings.position_embeddings.weight"]),
        spu.ptr(t["embeddings.token_type_embeddings.weight"]), spu.ptr(t["embeddings.LayerNorm.weight"]),
        spu.ptr(t["embeddings.LayerNorm.bias"]), spec.ln_eps, spec.vocab, spu.ptr(x32), spu.ptr(x16),
        spu.ptr(mb), B, S, H, spu.stream_ptr()))
    _sync()
    assert np.allclose(x32.cpu().numpy().reshape(B, S, H), layers[0], rtol=1e-4, atol=1e-4)
    assert torch.equal(x16.cpu(), x32.cpu().to(torch.bfloat16))
    assert np.array_equal(mb.cpu().numpy().reshape(B, S) == 0.0, mask != 0)
    assert np.all(mb.cpu().numpy().reshape(B, S)[mask == 0] == np.float32(-1e30))


# ------------------------------------------------- gather / scatter: bit-exact
def _dev_table(ptrs):
    return torch.tensor(ptrs, dtype=torch.int64).cuda()


@pytest.mark.parametrize("lens,seq", [([128] * 8, 128), ([16, 512, 77, 1, 300, 0, 511], 512),
                                      ([5], 8), ([3, 3, 3], 3)])
@pytest.mark.parametrize("with_mask", [False, True])
def test_gather_pad_mask_bit_exact(spu, oracle_gs, lens, seq, with_mask):
    rng = np.random.Generator(np.random.PCG64(len(lens) * 1000 + seq))
    n = len(lens)
    src = [rng.integers(-2**62, 2**62, size=max(L, 1), dtype=np.int64) for L in lens]
    msk = [rng.integers(0, 10, size=max(L, 1), dtype=np.int64) for L in lens]
    # oracle
    ids_o = np.full((n, seq), -1, np.int64)
    mask_o = np.full((n, seq), -1, np.int64)
    PP = C.c_void_p * n
    lens_np = np.asarray(lens, np.int32)
    rc = oracle_gs.oracle_gather_pad_mask(PP(*[s.ctypes.data for s in src]),
                                          PP(*[m.ctypes.data for m in msk]) if with_mask else None,
                                          lens_np.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.c_size_t(seq),
                                          C.c_int64(0), ids_o.ctypes.data_as(C.c_void_p),
                                          mask_o.ctypes.data_as(C.c_void_p))
    assert rc == 0
    # device
    dsrc = [torch.from_numpy(s).cuda() for s in src]
    dmsk = [torch.from_numpy(m).cuda() for m in msk]
    ids_d = torch.full((n, seq), -1, dtype=torch.int64, device="cuda")
    mask_d = torch.full((n, seq), -1, dtype=torch.int64, device="cuda")
    tab_i, tab_m = _dev_table([t.data_ptr() for t in dsrc]), _dev_table([t.data_ptr() for t in dmsk])
    lens_d = torch.from_numpy(lens_np).cuda()  # device temporaries must outlive the async call
    spu.check(spu.lib().spu_gather_pad_mask_async(
        spu.ptr(tab_i), spu.ptr(tab_m) if with_mask else None,
        spu.ptr(lens_d), n, seq, 0, spu.ptr(ids_d), spu.ptr(mask_d),
        spu.stream_ptr()))
    _sync()
    assert np.array_equal(ids_d.cpu().numpy(), ids_o)
    assert np.array_equal(mask_d.cpu().numpy(), mask_o)


def test_gather_full_length_equals_reference_concat(spu, oracle_gs):
    """All lens == seq: stage 1 must reproduce copy_job_inputs_to_slot byte for byte."""
    rng = np.random.Generator(np.random.PCG64(9))
    n, seq = 6, 128
    src = [rng.integers(0, 30522, size=seq, dtype=np.int64) for _ in range(n)]
    slot = np.zeros(n * seq, np.int64)
    nb = (C.c_size_t * n)(*[s.nbytes for s in src])
    ne = (C.c_size_t * n)(*[s.size for s in src])
    nx, by = C.c_size_t(), C.c_size_t()
    rc = oracle_gs.oracle_gather_concat((C.c_void_p * n)(*[s.ctypes.data for s in src]), nb, ne, C.c_size_t(n),
                                        slot.ctypes.data_as(C.c_void_p), C.c_size_t(slot.nbytes),
                                        C.byref(nx), C.byref(by))
    assert rc == 0 and nx.value == n * seq
    dsrc = [torch.from_numpy(s).cuda() for s in src]
    ids_d = torch.empty(n, seq, dtype=torch.int64, device="cuda")
    mask_d = torch.empty(n, seq, dtype=torch.int64, device="cuda")
    tab = _dev_table([t.data_ptr() for t in dsrc])
    lens_d = torch.full((n,), seq, dtype=torch.int32, device="cuda")
    spu.check(spu.lib().spu_gather_pad_mask_async(
        spu.ptr(tab), None, spu.ptr(lens_d), n, seq, 0, spu.ptr(ids_d),
        spu.ptr(mask_d), spu.stream_ptr()))
    _sync()
    assert np.array_equal(ids_d.cpu().numpy().ravel(), slot)
    assert (mask_d == 1).all()


@pytest.mark.parametrize("sizes", [[393216] * 5, [4000, 8000, 4000], [1, 17, 393216 + 3, 0, 64]])
def test_scatter_and_gather_segments_bit_exact(spu, oracle_gs, sizes):
    rng = np.random.Generator(np.random.PCG64(sum(sizes)))
    total = sum(sizes)
    agg = rng.integers(0, 256, size=total, dtype=np.uint8)
    offs = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.uint64)
    # oracle: contiguous dim-0 slices (rows are bytes here)
    ref = []
    for o, s in zip(offs, sizes):
        d = np.zeros(max(s, 1), np.uint8)
        got = C.c_int64()
        if s:
            rc = oracle_gs.oracle_scatter_slice(agg.ctypes.data_as(C.c_void_p), C.c_size_t(total), C.c_size_t(total),
                                                C.c_int64(int(o)), C.c_int64(s), d.ctypes.data_as(C.c_void_p),
                                                C.byref(got))
            assert rc == 0 and got.value == s
        ref.append(d[:s])
    agg_d = torch.from_numpy(agg).cuda()
    dsts = [torch.zeros(max(s, 1), dtype=torch.uint8, device="cuda") for s in sizes]
    off_d = torch.from_numpy(offs.astype(np.int64)).cuda()
    nb_d = torch.tensor(sizes, dtype=torch.int64).cuda()
    tab = _dev_table([t.data_ptr() for t in dsts])
    spu.check(spu.lib().spu_scatter_segments_async(spu.ptr(agg_d), spu.ptr(off_d), spu.ptr(nb_d),
                                                   spu.ptr(tab), len(sizes), spu.stream_ptr()))
    _sync()
    for d, r, s in zip(dsts, ref, sizes):
        assert np.array_equal(d.cpu().numpy()[:s], r)
    # and back: gather the per-request buffers into one contiguous batch (round trip)
    back = torch.zeros(max(total, 1), dtype=torch.uint8, device="cuda")
    spu.check(spu.lib().spu_gather_segments_async(spu.ptr(tab), spu.ptr(off_d), spu.ptr(nb_d),
                                                  len(sizes), spu.ptr(back), spu.stream_ptr()))
    _sync()
    assert np.array_equal(back.cpu().numpy()[:total], agg)


def test_empty_batches_are_noops(spu):
    assert spu.lib().spu_gather_pad_mask_async(None, None, None, 0, 128, 0, None, None, spu.stream_ptr()) == 0
    assert spu.lib().spu_scatter_segments_async(None, None, None, None, 0, spu.stream_ptr()) == 0
    assert spu.lib().spu_affine_f32_async(None, None, 0, 1.0, 0.0, spu.stream_ptr()) == 0


def test_affine_known_answers(spu):
    """x+1 on {1,2,3} -> {2,3,4} (reference tests/integration/starpu/integration_starpu_setup.cpp:42-60)
    and x+1.5 (tests/unit/core/unit_starpu_setup.cpp:2332-2433)."""
    x = torch.tensor([1.0, 2.0, 3.0], device="cuda")
    y = torch.empty_like(x)
    spu.check(spu.lib().spu_affine_f32_async(spu.ptr(x), spu.ptr(y), 3, 1.0, 1.0, spu.stream_ptr()))
    _sync()
    assert y.cpu().tolist() == [2.0, 3.0, 4.0]
    spu.check(spu.lib().spu_affine_f32_async(spu.ptr(x), spu.ptr(y), 3, 1.0, 1.5, spu.stream_ptr()))
    _sync()
    assert y.cpu().tolist() == [2.5, 3.5, 4.5]
