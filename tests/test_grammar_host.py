"""The grammar automaton (csrc/grammar.cpp) against the reference's (whisper.cpp:5484-5896) WITHOUT a device: same vocabulary on both
sides, random walks through the grammars of tests/test_grammar.py plus a few that stress other corners; at every step the reject set over
the whole vocabulary and the number of live stacks must be the reference's.  (tests/test_grammar.py repeats four of them through a
context on the GPU and checks whole transcripts.)"""
import ctypes as C

import numpy as np
import pytest

from test_grammar import ALT, CHAR, CHAR_ALT, CHAR_NOT, GRAMMARS, RNG_UPPER, RULE_REF, c, compile_rules
from whisper_rust_b200 import WhisperContext, WhisperContextParameters, synth

MORE = {
    # root ::= [一-鿿] root | [一-鿿]      (three-byte UTF-8: lead byte, then two continuation bytes arriving as separate tokens)
    "cjk": [[(CHAR, 0x4E00), (RNG_UPPER, 0x9FFF), (RULE_REF, 0), (ALT, 0), (CHAR, 0x4E00), (RNG_UPPER, 0x9FFF)]],
    # root ::= word (" " word)* ; written as root ::= word rest ; rest ::= " " word rest | <empty> ; word ::= [a-z] tail ; tail ::= [a-z] tail | <empty>
    "words": [[(RULE_REF, 2), (RULE_REF, 1)],
              [(CHAR, c(" ")), (RULE_REF, 2), (RULE_REF, 1), (ALT, 0)],
              [(CHAR, c("a")), (RNG_UPPER, c("z")), (RULE_REF, 3)],
              [(CHAR, c("a")), (RNG_UPPER, c("z")), (RULE_REF, 3), (ALT, 0)]],
    # root ::= ("yes" | "no" | "t" [0-9]) "." : three alternatives of different lengths sharing nothing, then a terminator: the grammar ends
    "choice": [[(RULE_REF, 1), (CHAR, c("."))],
               [(CHAR, c("y")), (CHAR, c("e")), (CHAR, c("s")), (ALT, 0), (CHAR, c("n")), (CHAR, c("o")), (ALT, 0), (CHAR, c("t")), (CHAR, c("0")),
                (RNG_UPPER, c("9"))]],
    # root ::= [^\x00-\x7f] root | [a-c] : a negated ASCII range (every multi-byte character passes), partial sequences against CHAR_NOT
    "nonascii": [[(CHAR_NOT, 0x00), (RNG_UPPER, 0x7F), (RULE_REF, 0), (ALT, 0), (CHAR, c("a")), (RNG_UPPER, c("c"))]],
    # root ::= [ a-e] [f-j,] root | [k-o]   (classes with several CHAR_ALT members and ranges mixed)
    "classes": [[(CHAR, c(" ")), (CHAR_ALT, c("a")), (RNG_UPPER, c("e")), (CHAR, c("f")), (RNG_UPPER, c("j")), (CHAR_ALT, c(",")), (RULE_REF, 0), (ALT, 0),
                 (CHAR, c("k")), (RNG_UPPER, c("o"))]],
}
ALL = {**GRAMMARS, **MORE}


@pytest.fixture(scope="module")
def sides(prod_lib, ref_lib):
    sig_tail = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_int, C.c_void_p]
    prod_lib.whisper_b200_test_grammar_vocab.restype = C.c_int
    prod_lib.whisper_b200_test_grammar_vocab.argtypes = [C.c_void_p, C.c_int, C.c_int] + sig_tail
    ref_lib.ref_probe_grammar.restype = C.c_int
    ref_lib.ref_probe_grammar.argtypes = [C.c_void_p] + sig_tail
    blob = synth.model_bytes("micro", synth.InitRecipe(), with_tensors=False)      # the vocabulary is all the automaton needs
    cr = WhisperContext.new_from_buffer_with_params(blob, WhisperContextParameters(use_gpu=False), lib=ref_lib)
    V, eot = cr.n_vocab(), cr.token_eot()
    ref_lib.whisper_token_to_str.restype = C.c_char_p
    ref_lib.whisper_token_to_str.argtypes = [C.c_void_p, C.c_int]
    texts = [ref_lib.whisper_token_to_str(cr.ptr, i) or b"" for i in range(V)]     # the reference's own id -> text table
    arr = (C.c_char_p * V)(*texts)

    def ref_step(ptrs, n_rules, accepted, mask):
        acc = (C.c_int32 * max(1, len(accepted)))(*accepted)
        return ref_lib.ref_probe_grammar(cr.ptr, ptrs, n_rules, 0, acc, len(accepted), mask.ctypes.data)

    def prod_step(ptrs, n_rules, accepted, mask):
        acc = (C.c_int32 * max(1, len(accepted)))(*accepted)
        return prod_lib.whisper_b200_test_grammar_vocab(arr, V, eot, ptrs, n_rules, 0, acc, len(accepted), mask.ctypes.data)

    yield ref_step, prod_step, texts, V, eot
    cr.free()


@pytest.mark.parametrize("name", list(ALL))
def test_reject_sets_equal_the_reference_on_the_host(sides, name):
    ref_step, prod_step, texts, V, eot = sides
    ptrs, keep = compile_rules(ALL[name])
    rng = np.random.default_rng(7)
    n_checked, n_rejecting = 0, 0
    for walk in range(8):
        accepted = []
        for step in range(14):
            ma, mb = np.zeros(V, dtype=np.uint8), np.zeros(V, dtype=np.uint8)
            na = ref_step(ptrs, len(keep), accepted, ma)
            nb = prod_step(ptrs, len(keep), accepted, mb)
            assert na == nb, (name, accepted)
            np.testing.assert_array_equal(ma, mb, err_msg=f"{name} after {accepted}")
            n_checked += 1
            n_rejecting += int(ma.any())
            ok = [i for i in range(eot) if ma[i] == 0 and texts[i]]
            if na == 0 or not ok:
                break
            # mostly tokens the grammar accepts; now and then one it rejects (the automaton then dies: zero stacks on both sides)
            pool = ok if rng.random() > 0.1 else [i for i in range(eot) if ma[i] == 1][:2000] or ok
            accepted.append(int(rng.choice(pool)))
    assert n_checked >= 16 and n_rejecting >= 8


def test_inactive_grammar_rejects_nothing(sides, prod_lib):
    """a start rule out of range gives no stacks (the reference asserts there, wc:5784-5822; the product treats it as "no grammar"):
    nothing is penalised"""
    ptrs, keep = compile_rules(ALL["fixed"])
    m = np.ones(2, dtype=np.uint8)
    acc = (C.c_int32 * 1)(0)
    n = prod_lib.whisper_b200_test_grammar_vocab((C.c_char_p * 2)(b"a", b"b"), 2, 2, ptrs, len(keep), len(keep), acc, 0, m.ctypes.data)
    assert n == 0 and not m.any()
