"""GPU: ``lgasr_row_logsumexp_noalloc`` (no counterpart in the reference)
against the existing ``lgasr_row_logsumexp``, through the C ABI.

The entry point allocates nothing: its scratch is a fixed file-scope ``__device__`` array inside the library whose device
address the library resolves with ``cudaGetSymbolAddress``.  Every call shares that array and finds what the last call
left there, so every shape is run TWICE IN A ROW (the second call must not be disturbed by the first one's leftovers),
and shapes of different split counts follow each other.  Tolerance: 2e-5 abs, the tolerance of the log-softmax tests
(``tests/test_gpu_device_cache.py``).  This file is named to be collected last.
"""

import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
ATOL = 2e-5
MAX_ROWS = 4096  # LGASR_NOALLOC_MAX_ROWS (include/lgasr.h)
MAX_VOCAB = 1 << 20  # LGASR_NOALLOC_MAX_VOCAB


def _call(name, logits, out, rows=None, vocab=None):
    from llm_guided_asr_b200 import _lib

    lib = _lib.load()
    r, v = logits.shape
    rows = r if rows is None else rows
    vocab = v if vocab is None else vocab
    with torch.cuda.device(logits.device):
        st = int(torch.cuda.current_stream(logits.device).cuda_stream)
        return getattr(lib, name)(C.c_void_p(logits.data_ptr()), 1 if logits.dtype == torch.bfloat16 else 0,
                                  C.c_int64(logits.stride(0)), C.c_void_p(out.data_ptr()), rows, vocab, C.c_void_p(st))


def _both(logits):
    from llm_guided_asr_b200 import _lib

    rows = logits.shape[0]
    ref = torch.full((rows,), float("nan"), dtype=torch.float32, device=logits.device)
    _lib.check(_call("lgasr_row_logsumexp", logits, ref))
    outs = []
    for _ in range(2):  # two calls in a row on the shared scratch
        out = torch.full((rows,), float("nan"), dtype=torch.float32, device=logits.device)
        _lib.check(_call("lgasr_row_logsumexp_noalloc", logits, out))
        outs.append(out)
    torch.cuda.synchronize()
    return ref, outs


@pytest.mark.parametrize(
    "rows,vocab,dtype",
    [
        (640, 5000, torch.float32),  # configs[1] rows: one slice per row
        (160, 50000, torch.float32),  # OWSM vocabulary: seven slices
        (64, 128256, torch.bfloat16),  # Llama-3 vocabulary, bf16 logits: eight slices
        (33, 32001, torch.float32),  # odd vocabulary: scalar loads, ragged last slice
        (MAX_ROWS, 1000, torch.float32),  # the row limit itself
        (3, 7, torch.bfloat16),  # tiny
    ],
)
def test_noalloc_matches_row_logsumexp_twice(rows, vocab, dtype):
    g = torch.Generator(device=DEV).manual_seed(rows * 131 + vocab)
    x = (4.0 * torch.randn(rows, vocab, device=DEV, generator=g)).to(dtype)
    x[0, : min(5, vocab)] = float("-inf")  # -inf entries travel through the (max, sum) merges
    ref, outs = _both(x)
    want = torch.logsumexp(x.float(), dim=-1)
    assert torch.isfinite(ref).all()
    assert (ref - want).abs().max().item() <= ATOL  # the existing entry point against torch (sanity of the checker)
    for out in outs:
        assert torch.isfinite(out).all()
        assert (out - ref).abs().max().item() <= ATOL
    assert torch.equal(outs[0], outs[1])  # slices are combined in slice order: the result does not depend on timing


def test_noalloc_strided_rows_and_all_minus_inf_row():
    """Rows inside a wider buffer (leading dimension > vocab, as the padded hand-off layout has) and a row without mass."""
    rows, vocab, ld = 50, 20000, 20480
    g = torch.Generator(device=DEV).manual_seed(9)
    buf = 3.0 * torch.randn(rows, ld, device=DEV, generator=g)
    x = buf[:, :vocab]
    x[7] = float("-inf")
    ref, outs = _both(x)
    for out in outs:
        assert out[7].item() == float("-inf")  # torch.logsumexp's answer for a row without mass
        keep = torch.arange(rows, device=DEV) != 7
        assert (out[keep] - ref[keep]).abs().max().item() <= ATOL


def test_noalloc_rejects_sizes_beyond_the_scratch():
    from llm_guided_asr_b200 import _lib

    lib = _lib.load()
    lib.lgasr_last_error.restype = C.c_char_p
    x = torch.zeros(8, 16, device=DEV)
    out = torch.zeros(8, device=DEV)
    assert _call("lgasr_row_logsumexp_noalloc", x, out, rows=MAX_ROWS + 1) != 0
    assert b"LGASR_NOALLOC_MAX_ROWS" in lib.lgasr_last_error()
    assert _call("lgasr_row_logsumexp_noalloc", x, out, vocab=MAX_VOCAB + 1) != 0  # also ld_in < vocab: refused either way
    big = torch.zeros(1, 8, device=DEV)
    rc = getattr(lib, "lgasr_row_logsumexp_noalloc")(C.c_void_p(big.data_ptr()), 0, C.c_int64(MAX_VOCAB + 8),
                                                      C.c_void_p(out.data_ptr()), 1, MAX_VOCAB + 8, C.c_void_p(0))
    assert rc != 0 and b"LGASR_NOALLOC_MAX_VOCAB" in lib.lgasr_last_error()
    # and a legal call right after the refusals still works
    ok = torch.full((8,), float("nan"), device=DEV)
    _lib.check(_call("lgasr_row_logsumexp_noalloc", x, ok))
    torch.cuda.synchronize()
    assert (ok - torch.log(torch.tensor(16.0))).abs().max().item() <= ATOL


def test_noalloc_call_allocates_no_device_memory():
    """The driver's free-memory figure does not move across calls (after one warm-up call that resolves the symbol)."""
    x = torch.randn(256, 50000, device=DEV)
    out = torch.empty(256, device=DEV)
    from llm_guided_asr_b200 import _lib

    _lib.check(_call("lgasr_row_logsumexp_noalloc", x, out))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(5):
        _lib.check(_call("lgasr_row_logsumexp_noalloc", x, out))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
