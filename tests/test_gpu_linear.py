"""b2p_linear (tcgen05/TMEM/TMA GEMM) through the C ABI against a float64 reference.

Layout bugs (swizzle / descriptor / TMEM lane mapping) give O(1) errors, so two kinds of
check are used: (a) small-integer operands, where every product and partial sum is
exactly representable -> results must be bit-exact in both operand kinds; (b) random
operands against float64 on the same (bf16-rounded) inputs with a tolerance that only
leaves room for fp32 accumulation order (bf16) or the tf32 input rounding (tf32)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _split(engine, x):
    hi, lo = torch.empty_like(x), torch.empty_like(x)
    rc = engine._lib.lib().b2p_split_tf32(C.c_void_p(x.data_ptr()), C.c_void_p(hi.data_ptr()),
                                          C.c_void_p(lo.data_ptr()), x.numel(), None)
    assert rc == 0
    return hi, lo


def _linear(engine, kind, a, w, bias=None, res=None, act=0, out_bf16=False, tile_n=0):
    """kind 0 bf16, 1 tf32, 2 tf32x3 (operands are split here); tile_n forces the N tile width."""
    M, K = a.shape
    N = w.shape[0]
    out = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16 if out_bf16 else torch.float32)
    p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
    stream = torch.cuda.current_stream(torch.device(DEV)).cuda_stream
    a_lo = w_lo = None
    if kind == 2:
        assert a.is_contiguous() and w.is_contiguous()
        a, a_lo = _split(engine, a)
        w, w_lo = _split(engine, w)
    rc = engine._lib.lib().b2p_linear(kind | (tile_n << 8), p(a), p(a_lo), a.stride(0), p(w), p(w_lo), w.stride(0),
                                      p(bias), p(res), res.stride(0) if res is not None else 0, p(out),
                                      out.stride(0), int(out_bf16), act, M, N, K, C.c_void_p(stream))
    assert rc == 0, engine._lib.lib().b2p_last_error()
    torch.cuda.synchronize()
    return out


def _ref(a, w, bias, res, act):
    y = a.double() @ w.double().t()
    if bias is not None:
        y = y + bias.double()
    if act == 1:
        y = torch.relu(y)
    elif act == 2:
        y = torch.nn.functional.gelu(y)
    if res is not None:
        y = y + res.double()
    return y


SHAPES = [(128, 128, 64), (128, 128, 512), (256, 384, 256), (200, 136, 72), (1, 10, 64), (130, 1001, 768),
          (2048, 3072, 768), (128, 512, 4096)]


# tile_n 2 = the CTA-pair kernel (tcgen05 cta_group::2, one 256 x 256 tile per pair of SMs)
@pytest.mark.parametrize("kind,tile_n", [(0, 128), (0, 256), (0, 2), (1, 128), (1, 256), (2, 128)])
@pytest.mark.parametrize("shape", SHAPES + [(1000, 1300, 200), (4096, 4096, 1024)])
def test_exact_on_small_integers(cuda_engine, kind, tile_n, shape):
    M, N, K = shape
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K)
    dt = torch.bfloat16 if kind == 0 else torch.float32
    a = torch.randint(-2, 3, (M, K), generator=g).to(dt).to(DEV)
    w = torch.randint(-2, 3, (N, K), generator=g).to(dt).to(DEV)
    bias = torch.randint(-3, 4, (N,), generator=g).float().to(DEV)
    res = torch.randint(-3, 4, (M, N), generator=g).float().to(DEV)
    out = _linear(cuda_engine, kind, a, w, bias, res, act=1, tile_n=tile_n)
    want = _ref(a, w, bias, res, 1).float()
    assert torch.equal(out, want), f"max diff {(out - want).abs().max().item()}"


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("act,use_res,out_bf16", [(0, False, False), (2, False, True), (0, True, False)])
@pytest.mark.parametrize("tile_n", [0, 2])
def test_bf16_random(cuda_engine, shape, act, use_res, out_bf16, tile_n):
    M, N, K = shape
    if out_bf16 and N % 2:
        pytest.skip("bf16 output needs even N")
    g = torch.Generator().manual_seed(K + N)
    a = torch.randn(M, K, generator=g).to(torch.bfloat16).to(DEV)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(torch.bfloat16).to(DEV)
    bias = torch.randn(N, generator=g).to(DEV)
    res = torch.randn(M, N, generator=g).to(DEV) if use_res else None
    out = _linear(cuda_engine, 0, a, w, bias, res, act, out_bf16, tile_n=tile_n)
    want = _ref(a, w, bias, res, act)
    tol = 2e-2 if out_bf16 else 1e-4
    assert (out.double() - want).abs().max().item() <= tol * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("shape", [(128, 128, 256), (200, 136, 72), (2048, 3072, 768)])
def test_tf32_random(cuda_engine, shape):
    M, N, K = shape
    g = torch.Generator().manual_seed(K)
    a = torch.randn(M, K, generator=g).to(DEV)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(DEV)
    out = _linear(cuda_engine, 1, a, w)
    want = _ref(a, w, None, None, 0)
    assert (out.double() - want).abs().max().item() <= 2e-2       # tf32 inputs: 2^-10 relative per product


@pytest.mark.parametrize("shape", [(128, 128, 256), (200, 136, 72), (2048, 3072, 768), (128, 512, 4096)])
def test_tf32x3_is_fp32_grade(cuda_engine, shape):
    """3xTF32: hi/lo operand planes, three MMAs per K step -> error at the level of an fp32 FMA chain."""
    M, N, K = shape
    g = torch.Generator().manual_seed(K + 1)
    a = torch.randn(M, K, generator=g).to(DEV)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(DEV)
    hi, lo = _split(cuda_engine, a)
    assert ((hi.view(torch.int32) & 0x1FFF) == 0).all() and ((lo.view(torch.int32) & 0x1FFF) == 0).all()
    assert ((a.double() - hi.double() - lo.double()).abs() <= a.abs().double() * 2.0 ** -21).all()
    out = _linear(cuda_engine, 2, a, w)
    want = _ref(a, w, None, None, 0)
    err = (out.double() - want).abs().max().item()
    fp32_err = ((a @ w.t()).double() - want).abs().max().item()      # what an fp32 GEMM leaves
    print(f"\ntf32x3 {shape}: max err {err:.2e} (torch fp32 matmul: {fp32_err:.2e})")
    # The products are fp32-grade; what remains is the tensor core's fp32 ACCUMULATOR, which
    # truncates instead of rounding, so the error grows linearly with the number of K steps
    # (measured ~7.5e-9 * K relative to the largest output).  The stages keep K per accumulator
    # short where it matters (fc1 runs as 16 K-slices summed in fp32 on the CUDA cores).
    assert err <= 2.5e-8 * K * want.abs().max().item()


def test_strided_operands_and_argument_errors(cuda_engine):
    L = cuda_engine._lib
    g = torch.Generator().manual_seed(1)
    big = torch.randn(256, 1024, generator=g).to(torch.bfloat16).to(DEV)
    a = big[:, 256:768]                                      # lda = 1024, offset 512 B
    w = torch.randn(128, 512, generator=g).to(torch.bfloat16).to(DEV)
    out = _linear(cuda_engine, 0, a, w)
    assert (out.double() - _ref(a, w, None, None, 0)).abs().max().item() <= 1e-3
    bad = big[:, 1:513]                                      # 2-byte offset: not 16-byte aligned
    o = torch.empty(256, 128, device=DEV)
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = L.lib().b2p_linear(0, p(bad), None, 1024, p(w), None, 512, None, None, 0, p(o), 128, 0, 0, 256, 128, 512, None)
    assert rc == L.EINVAL and b"16-byte" in L.lib().b2p_last_error()
    assert L.lib().b2p_linear(7, p(a), None, 1024, p(w), None, 512, None, None, 0, p(o), 128, 0, 0, 256, 128, 512,
                              None) == L.EINVAL
    assert L.lib().b2p_linear(2, p(a), None, 1024, p(w), None, 512, None, None, 0, p(o), 128, 0, 0, 256, 128, 512,
                              None) == L.EINVAL          # tf32x3 without lo planes


def test_throughput_report(cuda_engine):
    """Not an assertion on speed -- prints achieved TFLOP/s for the GPT shapes (pytest -s)."""
    shapes = [(4096, 4096, 4096), (8192, 8192, 8192), (2048, 3072, 1024), (2048, 1024, 4096), (2048, 50304, 1024),
              (2048, 2304, 768), (32768, 3072, 768)]
    for (M, N, K) in shapes:
        for kind, tile_n in [(0, 128), (0, 256), (0, 2), (1, 256), (2, 128)]:
            if kind != 0 and M * N * K > 4096 ** 3:
                continue
            dt = torch.bfloat16 if kind == 0 else torch.float32
            a = torch.randn(M, K, device=DEV).to(dt)
            w = torch.randn(N, K, device=DEV).to(dt)
            out = torch.empty(M, N, device=DEV)
            p = lambda t: C.c_void_p(t.data_ptr())
            s = torch.cuda.current_stream().cuda_stream
            run = lambda: cuda_engine._lib.lib().b2p_linear(kind | (tile_n << 8), p(a), p(a), K, p(w), p(w), K, None,
                                                            None, 0, p(out), N, 0, 0, M, N, K, C.c_void_p(s))
            for _ in range(3):
                assert run() == 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                run()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            name = {0: "bf16", 1: "tf32", 2: "tf32x3"}[kind]
            print(f"\nlinear {name:6s} BN={tile_n} {M}x{N}x{K}: {ms * 1e3:8.1f} us  {2 * M * N * K / ms / 1e9:7.1f} TFLOP/s")
