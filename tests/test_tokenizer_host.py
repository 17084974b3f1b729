"""The tokenizer (csrc/abi.cu: tokenize_text; reference whisper.cpp:3288-3336: GPT-2 style pre-split, then greedy longest match) against
the reference WITHOUT a device: the reference's own id -> text table on both sides, a corpus of hand-picked corner cases and a few
thousand random strings.  (tests/test_gpu_api_rows.py checks whisper_tokenize through a context on the GPU.)"""
import ctypes as C

import numpy as np
import pytest

from whisper_rust_b200 import WhisperContext, WhisperContextParameters, synth

CORNERS = [
    "", " ", "  ", "   ", "a", " a", "  a", "a ", "a  ", " a b ", "\t", "\n", "\n\n", " \n", "\n ", "\ta", "a\tb", "a \t b", "a\n\nb", "\r\n",
    "it's", "It's", "IT'S", "we're we've I'm they'll he'd don't", "'s", "'", "''", "'x", "x'", "'re'", "'llama", "a'sb", " 's", "'S", "' s",
    "t1 t12 t123 t1234 t50000 t50256 t50257 t99999", " t300 t300  t300", "t", "tt1", "t1t2", "t 1", "t-1", "T1", "t300t301 t302",
    "123", " 123", "1 2 3", "12ab34", "ab12cd", "1.5", "1,000,000", "3rd", "x2", "2x", " 2x ", "٣", "１２",
    "hello, world!", "hello ,world", "wait... what?!", "a.b", "a . b", "(a)", " (a) ", "--", " --", "-- ", "a--b", "@#$%^&*", " @ # ",
    "naïve café", "über", "日本語", " 日本語 ", "日本 語", "emoji 😀 ok", "😀😀", "a😀b", "é", " é", "é ", " x", "x ", " ",
    "mixed\ttabs\nand lines\r\nhere", "trailing space ", "trailing spaces   ", "   leading", "mid   dle", "end\n", "end \n", "end\n ",
    "[_EOT_]", "[_BEG_]", "<|endoftext|>", "[_TT_10]", "a[_EOT_]b", "\x01\x02", "a\x7fb", "<0>", "\x0b\x0c",
]


@pytest.fixture(scope="module")
def sides(prod_lib, ref_lib):
    prod_lib.whisper_b200_test_tokenize_vocab.restype = C.c_int
    prod_lib.whisper_b200_test_tokenize_vocab.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    ref_lib.whisper_tokenize.restype = C.c_int
    ref_lib.whisper_tokenize.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int]
    ref_lib.whisper_token_to_str.restype = C.c_char_p
    ref_lib.whisper_token_to_str.argtypes = [C.c_void_p, C.c_int]
    blob = synth.model_bytes("micro", synth.InitRecipe(), with_tensors=False)
    cr = WhisperContext.new_from_buffer_with_params(blob, WhisperContextParameters(use_gpu=False), lib=ref_lib)
    V = cr.n_vocab()
    texts = [ref_lib.whisper_token_to_str(cr.ptr, i) or b"" for i in range(V)]
    arr = (C.c_char_p * V)(*texts)

    def both(strings, n_max=256):
        """[(count, ids)] from the reference and from the product for a list of byte strings"""
        buf = (C.c_int32 * n_max)()
        ref = []
        for s in strings:
            n = ref_lib.whisper_tokenize(cr.ptr, s, buf, n_max)
            ref.append((n, list(buf[:max(n, 0)])))
        out = np.zeros((len(strings), n_max), dtype=np.int32)
        counts = np.zeros(len(strings), dtype=np.int32)
        assert prod_lib.whisper_b200_test_tokenize_vocab(arr, V, (C.c_char_p * len(strings))(*strings), len(strings), out.ctypes.data, n_max,
                                                         counts.ctypes.data) == 0
        return ref, [(int(c), out[k, :max(int(c), 0)].tolist()) for k, c in enumerate(counts)]

    yield both, texts
    cr.free()


def test_corner_cases_tokenize_like_the_reference(sides):
    both, texts = sides
    strings = [s.encode("utf-8") for s in CORNERS]
    ref, got = both(strings)
    for s, a, b in zip(strings, ref, got):
        assert a == b, repr(s)
        if a[0] > 0 and b"<0>" not in s:
            assert b"".join(texts[i] for i in a[1]) == s, repr(s)                      # and the pieces spell the input


def test_random_strings_tokenize_like_the_reference(sides):
    both, _ = sides
    rng = np.random.default_rng(5)
    pieces = list("abcxyzTtABZ") + list("0123456789") + [" ", " ", " ", "  ", "\t", "\n", "\r", "'", "'s", "'t", "'re", "'ve", "'m", "'ll", "'d",
              ".", ",", "!", "?", "-", "_", "(", ")", "\"", "é", "ß", "日", "語", "😀", " ", "t1", "t25", " t300", "t4096", "t50256", "[", "]", "_"]
    strings = ["".join(rng.choice(pieces, size=int(rng.integers(1, 24)))).encode("utf-8") for _ in range(8000)]
    ref, got = both(strings)
    for s, a, b in zip(strings, ref, got):
        assert a == b, repr(s)
    assert sum(a[0] for a in ref) > 60000


def test_short_buffer_convention(sides):
    both, _ = sides
    ref, got = both([b"hello world, it's t123"], n_max=3)
    assert ref[0][0] == got[0][0] < 0 and got[0][1] == []                              # -(tokens needed), nothing written


def test_random_bytes_tokenize_like_the_reference(sides):
    """every byte value 1..255 (control characters, lone continuation bytes): the character classes are the "C" locale's on bytes"""
    both, _ = sides
    rng = np.random.default_rng(6)
    strings = [bytes(rng.integers(1, 256, size=int(rng.integers(1, 40))).astype(np.uint8)) for _ in range(3000)]
    strings += [bytes([b]) * k for b in range(1, 256) for k in (1, 2)]
    strings += [bytes([32, b]) for b in range(1, 256)] + [bytes([b, 32]) for b in range(1, 256)] + [bytes([39, b]) for b in range(1, 256)]
    ref, got = both(strings)
    for s, a, b in zip(strings, ref, got):
        assert a == b, repr(s)
