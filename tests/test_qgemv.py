"""oracle/qgemv.c (the restatement of llama.cpp's quantised CPU math that bench.py times as the closer CPU baseline)
against gguf-py's reference de-quantiser: random Q4_K / Q4_0 blocks -> float weights by gguf-py; y = W x by the C code
(quantise x to Q8_K / Q8_0, integer block dot products) must equal W_deq @ dequant(x_q8) to fp32 rounding, scalar and
AVX-512 VNNI paths alike, and be within the activation-quantisation error of W_deq @ x."""
import numpy as np
import pytest

from oracle import qgemv

gguf = pytest.importorskip("gguf")
from gguf import GGMLQuantizationType as T   # noqa: E402
from gguf import quants                      # noqa: E402

pytestmark = pytest.mark.skipif(not qgemv.available(), reason="oracle/liboracle_qgemv.so not built (make -C oracle)")

FMT = {"q4_K": (T.Q4_K, 144, 256), "q4_0": (T.Q4_0, 18, 32)}


def _blocks(fmt, n, k, rng):
    _, bb, per = FMT[fmt]
    nb = n * k // per
    raw = rng.integers(0, 256, (nb, bb), dtype=np.uint8)
    raw[:, 0:2] = np.abs(rng.normal(0, 0.01, nb)).astype(np.float16).view(np.uint8).reshape(nb, 2)
    if fmt == "q4_K":
        raw[:, 2:4] = np.abs(rng.normal(0, 0.005, nb)).astype(np.float16).view(np.uint8).reshape(nb, 2)
    return raw


def _dequant_x(fmt, xq, k):
    if fmt == "q4_0":
        return quants.dequantize(xq.reshape(1, -1), T.Q8_0).reshape(-1).astype(np.float64)
    blk = xq.reshape(k // 256, 292)
    d = blk[:, 0:4].copy().view(np.float32).reshape(-1, 1).astype(np.float64)
    q = blk[:, 4:260].copy().view(np.int8).astype(np.float64)
    bs = blk[:, 260:292].copy().view(np.int16).reshape(-1, 16)
    assert (q.reshape(-1, 16, 16).sum(-1) == bs).all()          # the sums of 16 the min term uses
    return (d * q).reshape(-1)


def test_block_sizes_match_ggml():
    assert qgemv.LIB.qgemv_sizeof_q4_K() == 144 and qgemv.LIB.qgemv_sizeof_q8_K() == 292
    assert qgemv.LIB.qgemv_sizeof_q4_0() == 18 and qgemv.LIB.qgemv_sizeof_q8_0() == 34
    assert 144 == quants.GGML_QUANT_SIZES[T.Q4_K][1] and 18 == quants.GGML_QUANT_SIZES[T.Q4_0][1]


@pytest.mark.parametrize("fmt", ["q4_K", "q4_0"])
@pytest.mark.parametrize("n,k", [(1, 256), (7, 512), (96, 2048)])
def test_gemv_matches_gguf_dequant(fmt, n, k):
    rng = np.random.default_rng(n * 1000 + k)
    raw = _blocks(fmt, n, k, rng)
    W = quants.dequantize(raw.reshape(n, -1), FMT[fmt][0]).astype(np.float64)
    x = rng.normal(0, 1, k).astype(np.float32)
    x[3] = 0.0
    xq = qgemv.quantize_activations(fmt, x)
    xd = _dequant_x(fmt, xq, k)
    assert np.abs(xd - x).max() <= np.abs(x).max() / 127 * 0.51 + 1e-6      # Q8 rounding: half a step of the block's scale
    want = W @ xd
    scale = np.abs(want).max() + 1e-9
    ys = qgemv.gemv(fmt, raw.reshape(-1), xq, n, k, scalar=True)
    yv = qgemv.gemv(fmt, raw.reshape(-1), xq, n, k)
    assert np.abs(ys - want).max() / scale <= 2e-5
    assert np.abs(yv - want).max() / scale <= 2e-5
    assert np.abs(yv - W @ x.astype(np.float64)).max() / scale <= 3e-2       # what quantising the activations costs


def test_zero_activations_and_q8_0_against_gguf_quantizer():
    x = np.zeros(512, np.float32)
    assert not qgemv.quantize_activations("q4_K", x).any()
    rng = np.random.default_rng(5)
    x = rng.normal(0, 2, 1024).astype(np.float32)
    mine = qgemv.quantize_activations("q4_0", x)
    ref = quants.quantize(x.reshape(1, -1), T.Q8_0).reshape(-1)
    assert np.array_equal(mine, ref)                                         # same blocks, bit for bit, as gguf-py's Q8_0


def test_decode_bench_runs_on_the_tiny_geometry():
    from oracle import llama_oracle as lo
    tok_s, secs, threads, nbytes = qgemv.decode_tok_s(lo.GEOMETRIES["small"], "q4_K", 3, 1)
    assert tok_s > 0 and threads >= 1 and nbytes > 0
