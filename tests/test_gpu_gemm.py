"""-m gpu parity tests for the tcgen05 GEMM (K5) through the C ABI: against the numpy oracle (linear.rs contract
y = x . W^T, bf16 in, f32 accumulate) at sizes the oracle finishes in seconds, and against the on-device
CUDA-core check kernel at the full Qwen3 shapes."""
import numpy as np
import pytest
import torch

from oracle import layers as O
from tests.util import assert_close, randn_bf16, to_dev_bf16, to_np_f32

pytestmark = pytest.mark.gpu

# f32 accumulation in a different order than numpy + one bf16 rounding of the output: 1 bf16 ulp (2^-8) + abs slack
ATOL, RTOL = 2e-2, 2 ** -7

SHAPES = [
    (128, 128, 64), (1, 8, 64), (5, 40, 128), (129, 136, 192), (256, 4096, 1024), (300, 1024, 2048),
    (256, 6144, 1024), (64, 1024, 3072), (7, 264, 4096), (512, 512, 512),
    (512, 768, 4096), (256, 1024, 3072), (130, 256, 2048),   # x-multicast clusters (narrow N, long K: TP shards)
]


@pytest.mark.parametrize("m,n,k", SHAPES)
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_gemm_vs_oracle(cuda, m, n, k, out_dtype):
    from nano_vllm_rs_b200 import ops
    rng = np.random.default_rng(m * 31 + n * 7 + k)
    x = randn_bf16(rng, (m, k)); w = randn_bf16(rng, (n, k), 0.05)
    y = ops.gemm_bf16(to_dev_bf16(x, cuda), to_dev_bf16(w, cuda), out_dtype=out_dtype)
    torch.cuda.synchronize()
    ref = O.linear(x, w, out_bf16=(out_dtype == torch.bfloat16))
    assert_close(to_np_f32(y), ref, ATOL if out_dtype == torch.bfloat16 else 2e-3, RTOL if out_dtype == torch.bfloat16 else 1e-4, f"gemm {m}x{n}x{k}")


@pytest.mark.parametrize("bn", ["32", "64", "128", "192", "256"])
def test_gemm_every_tile_width(cuda, bn, monkeypatch):
    """Force each BLOCK_N instantiation (the heuristic would otherwise hide some) in a fresh process."""
    import subprocess, sys, os
    code = (
        "import numpy as np, torch, sys; sys.path.insert(0, '.');"
        "from nano_vllm_rs_b200 import ops;"
        "g=torch.Generator(device='cuda').manual_seed(1);"
        "x=torch.randn((300,1024),generator=g,device='cuda').to(torch.bfloat16);"
        "w=(torch.randn((1160,1024),generator=g,device='cuda')*0.05).to(torch.bfloat16);"
        "y=ops.gemm_bf16(x,w,out_dtype=torch.float32); r=ops.gemm_bf16(x,w,out_dtype=torch.float32,simt=True);"
        "torch.cuda.synchronize(); d=(y-r).abs().max().item(); print('maxdiff',d); assert d<2e-3, d"
    )
    env = dict(os.environ, NVLLM_GEMM_BN=bn)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr


@pytest.mark.parametrize("m,n,k", [(256, 151936, 1024), (2048, 6144, 1024), (512, 24576, 4096), (512, 4096, 12288), (4096, 4096, 4096)])
def test_gemm_full_size_vs_device_check(cuda, m, n, k):
    """Full Qwen3 shapes (lm_head, prefill qkv, 8B gate_up / down): tcgen05 vs the CUDA-core check kernel (both
    accumulate in f32; only the summation order differs)."""
    from nano_vllm_rs_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    x = torch.randn((m, k), generator=g, device=cuda).to(torch.bfloat16)
    w = (torch.randn((n, k), generator=g, device=cuda) * 0.02).to(torch.bfloat16)
    y = ops.gemm_bf16(x, w, out_dtype=torch.float32)
    r = ops.gemm_bf16(x, w, out_dtype=torch.float32, simt=True)
    torch.cuda.synchronize()
    assert_close(to_np_f32(y), to_np_f32(r), 3e-3, 1e-4, f"gemm full {m}x{n}x{k}")
    # linearity (size-independent property): (2x) . W^T == 2 (x . W^T) exactly in f32 (power-of-two scale)
    y2 = ops.gemm_bf16((x.float() * 2).to(torch.bfloat16), w, out_dtype=torch.float32)
    torch.cuda.synchronize()
    assert torch.equal(y2, y * 2)


def test_gemm_strided_views(cuda):
    """Leading dimensions larger than the logical width (column slices of packed buffers, as o_proj reading the
    attention output or TP shards of a full weight)."""
    from nano_vllm_rs_b200 import ops
    rng = np.random.default_rng(3)
    xb = randn_bf16(rng, (70, 512)); wb = randn_bf16(rng, (200, 512), 0.05)
    xd, wd = to_dev_bf16(xb, cuda), to_dev_bf16(wb, cuda)
    out = torch.zeros((70, 400), dtype=torch.bfloat16, device=cuda)
    ops.gemm_bf16(xd[:, 128:384], wd[:, 256:512], out=out[:, 200:400])
    torch.cuda.synchronize()
    ref = O.linear(xb[:, 128:384], wb[:, 256:512])
    assert_close(to_np_f32(out[:, 200:400]), ref, ATOL, RTOL, "strided gemm")
    assert float(out[:, :200].abs().max()) == 0.0


@pytest.mark.parametrize("m,inter,k", [(256, 3072, 1024), (300, 1536, 512), (512, 12288, 4096), (64, 3072, 1024), (129, 256, 128)])
def test_gemm_silu_mul_fused_equals_unfused(cuda, m, inter, k, monkeypatch):
    """gate_up GEMM with the SiLU*mul epilogue (M > 128: CTA-pair kernel) is bit-identical to GEMM + silu_mul
    (split-K off: it changes the summation order, and is compared against the oracle in its own tests)."""
    from nano_vllm_rs_b200 import ops
    monkeypatch.setenv("NVLLM_GEMM_SPLITK", "0")
    g = torch.Generator(device="cuda").manual_seed(m + inter + k)
    x = torch.randn((m, k), generator=g, device=cuda).to(torch.bfloat16)
    w = (torch.randn((2 * inter, k), generator=g, device=cuda) * 0.05).to(torch.bfloat16)
    fused = ops.gemm_silu_mul(x, w)
    ref = ops.silu_mul(ops.gemm_bf16(x, w))
    torch.cuda.synchronize()
    assert torch.equal(fused, ref)
    if m * inter * k < 2 ** 28:   # and against the CPU oracle where it is cheap
        want = O.silu_and_mul(O.linear(to_np_f32(x), to_np_f32(w)))
        assert_close(to_np_f32(fused), want, ATOL, 2 ** -6, "gemm_silu_mul vs oracle")


@pytest.mark.parametrize("m,n,k,silu", [(512, 6144, 4096, False), (512, 6144, 4096, True), (256, 12288, 1024, True), (300, 1000, 512, False)])
def test_gemm_192_wide_tiles(cuda, m, n, k, silu, monkeypatch):
    """The 192-column tile (M = 512 TP decode batches on the wide projections: 4 x N/256 tiles fill 2/3 of the SMs or
    spill into a second wave) forced on plain and SiLU shapes, incl. an N that 192 does not divide: bit-identical to the
    256-wide tile (same k order, same f32 accumulation) and within tolerance of the CUDA-core check kernel."""
    from nano_vllm_rs_b200 import ops
    monkeypatch.setenv("NVLLM_GEMM_SPLITK", "0")
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    x = torch.randn((m, k), generator=g, device=cuda).to(torch.bfloat16)
    w = (torch.randn(((2 if silu else 1) * n, k), generator=g, device=cuda) * 0.03).to(torch.bfloat16)
    fn = (lambda: ops.gemm_silu_mul(x, w)) if silu else (lambda: ops.gemm_bf16(x, w))
    monkeypatch.setenv("NVLLM_GEMM_BN", "256")
    y256 = fn()
    monkeypatch.setenv("NVLLM_GEMM_BN", "192")
    y192 = fn()
    torch.cuda.synchronize()
    assert torch.equal(y192, y256)
    if not silu:
        r = ops.gemm_bf16(x, w, out_dtype=torch.float32, simt=True)
        assert_close(to_np_f32(y192), to_np_f32(r), ATOL, RTOL, "192-wide tile vs check kernel")


@pytest.mark.parametrize("kind", ["reference", "head_major"])
@pytest.mark.parametrize("m,H,Hkv,k,qknorm", [(256, 16, 8, 1024, True), (37, 4, 1, 512, True), (130, 8, 2, 256, False)])
def test_gemm_qkv_rope_store_fused_equals_unfused(cuda, kind, m, H, Hkv, k, qknorm, monkeypatch):
    """qkv GEMM with the q/k-norm + RoPE + KV-store epilogue vs nvllm_gemm_bf16 followed by nvllm_qknorm_rope_store
    (same rounding points; split-K off so that both GEMMs sum in the same order), for both cache layouts."""
    from nano_vllm_rs_b200 import ops
    monkeypatch.setenv("NVLLM_GEMM_SPLITK", "0")
    from nano_vllm_rs_b200._ffi import KvLayout
    D, bs = 128, 16
    g = torch.Generator(device="cuda").manual_seed(m + H + k)
    x = torch.randn((m, k), generator=g, device=cuda).to(torch.bfloat16)
    w = (torch.randn(((H + 2 * Hkv) * D, k), generator=g, device=cuda) * 0.05).to(torch.bfloat16)
    qw = (1 + 0.1 * torch.randn(D, generator=g, device=cuda)).to(torch.bfloat16) if qknorm else None
    kw = (1 + 0.1 * torch.randn(D, generator=g, device=cuda)).to(torch.bfloat16) if qknorm else None
    pos = torch.randint(0, 2048, (m,), generator=torch.Generator().manual_seed(1), dtype=torch.int64).to(cuda)
    nblk = (m + bs - 1) // bs + 2
    slots = torch.randperm(nblk * bs, generator=torch.Generator().manual_seed(2))[:m].to(torch.int32)
    if m > 4:
        slots[2] = -1
    slots = slots.to(cuda)
    table = ops.rope_table(2048, D, 1e6)
    lay = KvLayout.reference(bs, Hkv, D) if kind == "reference" else KvLayout.head_major(bs, Hkv, D)
    shape = (nblk, bs, Hkv, D) if kind == "reference" else (nblk, Hkv, bs, D)
    kc1, vc1 = torch.zeros(shape, dtype=torch.bfloat16, device=cuda), torch.zeros(shape, dtype=torch.bfloat16, device=cuda)
    kc2, vc2 = torch.zeros_like(kc1), torch.zeros_like(vc1)
    fused = ops.gemm_qkv_rope_store(x, w, pos, table, qw, kw, 1e-6, H, Hkv, D, kc1, vc1, slots, lay)
    ref = ops.gemm_bf16(x, w)
    ops.qknorm_rope_store(ref, pos, table, qw, kw, 1e-6, H, Hkv, D, kc2, vc2, slots, lay)
    torch.cuda.synchronize()
    # same rounding points; only the order of the 128-term sum of squares differs (thread-sequential vs warp shuffle), which
    # can move a normalised value by one bf16 ulp before the rotation
    assert_close(to_np_f32(fused), to_np_f32(ref), 2e-3, 2 ** -6, "fused qkv epilogue vs gemm + qknorm_rope_store")
    assert torch.equal(fused[:, (H + Hkv) * D:], ref[:, (H + Hkv) * D:])            # v heads: pure GEMM output
    assert (fused != ref).float().mean().item() < 0.02
    # the cache holds exactly the k / v heads the kernel wrote to the qkv buffer (bit-exact data movement, slot -1 skipped)
    kc3, vc3 = torch.zeros_like(kc1), torch.zeros_like(vc1)
    ops.store_kvcache(fused[:, H * D:(H + Hkv) * D].view(m, Hkv, D), fused[:, (H + Hkv) * D:].view(m, Hkv, D), kc3, vc3, slots, lay)
    torch.cuda.synchronize()
    assert torch.equal(kc1, kc3) and torch.equal(vc1, vc3)


SPLITK_CFGS = ["64,2", "64,4", "128,2", "128,4", "128,8", "256,2", "256,4", "256,8"]


@pytest.mark.parametrize("cfg", SPLITK_CFGS)
def test_gemm_splitk_every_config_vs_device_check(cuda, cfg, monkeypatch):
    """Cluster split-K GEMM (gemm_splitk.cu), every (BN, SK) instantiation forced through NVLLM_GEMM_SPLITK (re-read per
    call): M tail rows, K split 2/4/8 ways, against the CUDA-core check kernel and the same GEMM without split-K."""
    from nano_vllm_rs_b200 import ops
    m, n, k = (200, 1024, 2048) if cfg != "256,2" else (500, 2048, 2048)   # one case with four 128-row tile rows
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn((m, k), generator=g, device=cuda).to(torch.bfloat16)
    w = (torch.randn((n, k), generator=g, device=cuda) * 0.05).to(torch.bfloat16)
    r = ops.gemm_bf16(x, w, out_dtype=torch.float32, simt=True)
    monkeypatch.setenv("NVLLM_GEMM_SPLITK", "0")
    base = ops.gemm_bf16(x, w)
    monkeypatch.setenv("NVLLM_GEMM_SPLITK", cfg)
    y = ops.gemm_bf16(x, w)
    y2 = ops.gemm_bf16(x, w)
    torch.cuda.synchronize()
    assert_close(to_np_f32(y), to_np_f32(r), ATOL, RTOL, f"split-K {cfg} vs check kernel")
    assert torch.equal(y, y2), "fixed-order cluster reduction must be deterministic"
    # the un-split kernel sums the same products in another order: at most one bf16 ulp apart, mostly identical
    assert_close(to_np_f32(y), to_np_f32(base), 1e-3, 2 ** -7, f"split-K {cfg} vs un-split")
    assert (y != base).float().mean().item() < 0.05


@pytest.mark.parametrize("cfg", SPLITK_CFGS)
def test_gemm_splitk_silu_every_config_vs_oracle(cuda, cfg, monkeypatch):
    from nano_vllm_rs_b200 import ops
    m, inter, k = 200, 512, 1024
    rng = np.random.default_rng(5)
    x = randn_bf16(rng, (m, k)); w = randn_bf16(rng, (2 * inter, k), 0.05)
    xd, wd = to_dev_bf16(x, cuda), to_dev_bf16(w, cuda)
    monkeypatch.setenv("NVLLM_GEMM_SPLITK", cfg)
    y = ops.gemm_silu_mul(xd, wd)
    torch.cuda.synchronize()
    want = O.silu_and_mul(O.linear(x, w))
    assert_close(to_np_f32(y), want, ATOL, 2 ** -6, f"split-K silu {cfg} vs oracle")
    monkeypatch.setenv("NVLLM_GEMM_SPLITK", "0")
    base = ops.gemm_silu_mul(xd, wd)
    torch.cuda.synchronize()
    assert (y != base).float().mean().item() < 0.05


@pytest.mark.parametrize("m,n,k", [(256, 1024, 3072), (256, 1024, 2048), (4, 1024, 3072), (64, 2048, 4096), (1, 1024, 2048),
                                   (512, 768, 4096), (512, 4096, 6144), (300, 4096, 2048)])
def test_gemm_splitk_auto_decode_shapes(cuda, m, n, k):
    """The shapes the cost model sends to split-K by default (Qwen3-0.6B o_proj / down_proj at decode) vs the check kernel."""
    from nano_vllm_rs_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    x = torch.randn((m, k), generator=g, device=cuda).to(torch.bfloat16)
    w = (torch.randn((n, k), generator=g, device=cuda) * 0.02).to(torch.bfloat16)
    y = ops.gemm_bf16(x, w)
    r = ops.gemm_bf16(x, w, out_dtype=torch.float32, simt=True)
    torch.cuda.synchronize()
    assert_close(to_np_f32(y), to_np_f32(r), ATOL, RTOL, f"auto split-K {m}x{n}x{k}")


@pytest.mark.parametrize("m,n,k", [(512, 768, 4096), (256, 1024, 3072), (130, 256, 2048)])
def test_gemm_x_multicast_clusters_opt_in(cuda, m, n, k, monkeypatch):
    """The x-multicast cluster mode (off by default since the in-graph measurements) still computes the same GEMM."""
    from nano_vllm_rs_b200 import ops
    monkeypatch.setenv("NVLLM_GEMM_SPLITK", "0")
    monkeypatch.setenv("NVLLM_GEMM_MCAST", "1")
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    x = torch.randn((m, k), generator=g, device=cuda).to(torch.bfloat16)
    w = (torch.randn((n, k), generator=g, device=cuda) * 0.02).to(torch.bfloat16)
    y = ops.gemm_bf16(x, w)
    r = ops.gemm_bf16(x, w, out_dtype=torch.float32, simt=True)
    torch.cuda.synchronize()
    assert_close(to_np_f32(y), to_np_f32(r), ATOL, RTOL, f"mcast gemm {m}x{n}x{k}")
