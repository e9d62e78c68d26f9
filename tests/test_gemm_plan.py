"""GEMM kernel-selection policy (`b200_gemm_plan`): host logic of csrc/gemm.cu, no GPU needed.

The measurements behind the policy are in profiles/gemm_short_k_study.md and
profiles/r1_gpt2_gemm_shapes_final.txt; these tests pin what each class of problem runs on.
"""
import pytest

from b200ddp._native import bindings as B

T, C = 8192, 768   # GPT-2 small: tokens per rank, width


@pytest.mark.parametrize("M,N,K,kw", [
    (T, 3 * C, C, dict(bias=True)),                                   # qkv forward
    (T, C, C, dict(bias=True, aux=True)),                             # attention projection + residual
    (T, 4 * C, C, dict(bias=True, epilogue=B.EPI_GELU, aux_out=True)),  # fc + GELU + saved pre-activation
    (T, C, 4 * C, dict(bias=True, aux=True)),                         # fc2 + residual
    (T, 4 * C, C, dict(epilogue=B.EPI_GELU_BWD, aux=True, b_mn_major=True)),  # fc2 dgrad x GELU'
    (T, C, 3 * C, dict(b_mn_major=True)),                             # qkv dgrad
    (T, 50257, C, dict(ldd=50264)),                                   # LM head (ragged N, padded rows)
])
def test_transformer_forward_and_dgrad_run_on_sm_pairs_with_256_wide_tiles(M, N, K, kw):
    assert B.gemm_plan(M, N, K, **kw) == ("pair", 256)


@pytest.mark.parametrize("M,N", [(3 * C, C), (C, C), (4 * C, C), (C, 4 * C)])
def test_weight_gradients_take_256x128_pair_tiles(M, N):
    # few output tiles, K = 8192: the narrower pair tile fills more of the 74 SM pairs
    assert B.gemm_plan(M, N, T, b_mn_major=True) == ("pair", 128)


def test_long_reductions_with_a_full_wave_take_the_dual_accumulator_kernel():
    assert B.gemm_plan(8192, 8192, 8192) == ("m256", 256)
    assert B.gemm_plan(4096, 4096, 4096) == ("m256", 256)       # 256 tiles of 256x256 >= 148 SMs, K >= 4096
    assert B.gemm_plan(8192, 8192, 2048) == ("pair", 256)       # short K: double-buffered accumulators matter
    assert B.gemm_plan(T, C, 50257, b_mn_major=True) == ("pair", 256)   # LM-head dgrad: 96 tiles < 148


@pytest.mark.parametrize("kw", [dict(d_fp32=True), dict(colsum=True), dict(epilogue=B.EPI_RELU_DROPOUT, bias=True),
                                dict(epilogue=B.EPI_GATE_POSITIVE, aux=True), dict(ldd=2052)])
def test_heavy_epilogues_and_unaligned_outputs_stay_on_the_general_kernels(kw):
    assert B.gemm_plan(T, 2048, C, **kw)[0] == "general"


def test_fused_cross_entropy_tiles():
    assert B.gemm_plan(64, 10, 256, epilogue=B.EPI_SOFTMAX_CE, bias=True) == ("general", 16)
    assert B.gemm_plan(64, 32, 256, epilogue=B.EPI_SOFTMAX_CE) == ("general", 32)
    with pytest.raises(RuntimeError, match="N <= 32"):
        B.gemm_plan(64, 100, 256, epilogue=B.EPI_SOFTMAX_CE)


def test_split_k_takes_the_widest_tile_the_output_allows():
    assert B.gemm_plan(64, 576, 200704, d_fp32=True, split_k=1, b_mn_major=True) == ("general", 256)
    assert B.gemm_plan(64, 128, 200704, d_fp32=True, split_k=1, b_mn_major=True) == ("general", 128)
    assert B.gemm_plan(256, 64, 12544, d_fp32=True, split_k=1, b_mn_major=True) == ("general", 64)


def test_small_and_batched_problems():
    assert B.gemm_plan(64, 256, 784, bias=True)[0] == "lean"                  # one M tile: no pair
    assert B.gemm_plan(T, 64, 256)[0] == "lean"                               # N <= 64: no pair
    assert B.gemm_plan(1024, 1024, 64, batch=96, causal=1)[0] == "lean"       # attention-shaped: batched


def test_forced_choices_and_their_errors():
    assert B.gemm_plan(T, 2304, C, tile_n=128, cta_pair=1) == ("pair", 128)
    assert B.gemm_plan(T, 2304, C, tile_n=256, tile_m=256) == ("m256", 256)
    assert B.gemm_plan(T, 2304, C, tile_n=256, tile_m=128) == ("lean", 256)   # explicit tile: single-SM kernel
    assert B.gemm_plan(T, 2304, C, cta_pair=-1)[0] == "lean"
    with pytest.raises(RuntimeError, match="cta_pair"):
        B.gemm_plan(T, 2304, C, cta_pair=1, d_fp32=True)
    with pytest.raises(RuntimeError, match="tile_m 256"):
        B.gemm_plan(T, 2304, C, tile_n=256, tile_m=256, d_fp32=True)
    with pytest.raises(RuntimeError, match="tile_m must be"):
        B.gemm_plan(T, 2304, C, tile_m=64)


def test_weight_gradient_plan():
    from b200ddp.ops.functional import wgrad_plan
    # GPT-2 linears at 8192 tokens: enough pair tiles, written straight to the bf16 arena
    for n_out, n_in in ((3 * C, C), (4 * C, C), (C, 4 * C), (50257, C)):
        assert wgrad_plan(n_out, n_in, T, True) == "direct"
    # ResNet-50 stem / layer1 convolutions: tiny outputs, reductions over 200k-800k pixels
    assert wgrad_plan(64, 576, 64 * 56 * 56, True) == "split"
    assert wgrad_plan(64, 392, 64 * 112 * 112, True) == "split"
    assert wgrad_plan(256, 2304, 64 * 14 * 14, True) == "split"
    # fp32 arenas (fp32 wire dtype) cannot take the bf16 pair kernel
    assert wgrad_plan(3 * C, C, T, False) == "split"
    assert wgrad_plan(4096, 4096, 1024, False) == "plain"
    # SimpleNet-sized: short reduction, nothing to split
    assert wgrad_plan(256, 784, 64, True) == "direct" and wgrad_plan(256, 784, 64, False) == "plain"


from hypothesis import given, settings, strategies as st


@settings(max_examples=300, deadline=None)
@given(M=st.integers(1, 20000), N=st.integers(1, 60000), K=st.integers(1, 20000),
       d_fp32=st.booleans(), bias=st.booleans(), aux=st.booleans(), colsum=st.booleans(),
       epilogue=st.sampled_from([B.EPI_NONE, B.EPI_RELU_DROPOUT, B.EPI_GELU]), split=st.booleans(),
       b_mn=st.booleans(), sms=st.sampled_from([148, 132, 64]))
def test_auto_selection_always_lands_on_a_kernel_that_supports_the_problem(M, N, K, d_fp32, bias, aux, colsum,
                                                                           epilogue, split, b_mn, sms):
    """Whatever the problem, the automatic choice is a supported tile, and the large-tile kernels
    are only chosen for what they implement (bf16 output, light epilogue, no column sums, no
    split-K, more than one M tile)."""
    if split:   # split-K contract: plain fp32 output
        d_fp32, bias, aux, colsum, epilogue = True, False, False, False, B.EPI_NONE
    if b_mn and N < 64:
        N = 64      # MN-major B needs a 64-wide tile
    ldd = (N + 7) // 8 * 8
    kernel, bn = B.gemm_plan(M, N, K, sms=sms, d_fp32=d_fp32, bias=bias, aux=aux and N % 8 == 0, colsum=colsum,
                             epilogue=epilogue, split_k=int(split), b_mn_major=b_mn, ldd=ldd)
    assert bn in (16, 32, 64, 128, 256)
    if b_mn:
        assert bn >= 64
    if kernel in ("pair", "m256", "lean"):
        assert not d_fp32 and not colsum and not split and epilogue != B.EPI_RELU_DROPOUT and bn >= 64
    if kernel == "pair":
        assert M > 128 and N > 64 and bn in (128, 256)
    if kernel == "m256":
        assert bn == 256 and M >= 512 and K >= 4096
