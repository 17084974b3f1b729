// Grammar-constrained sampling (whisper.cpp:5484-5896; wh:161-189 whisper_grammar_element).  Host logic: a pushdown automaton over the
// UTF-8 code points of the token texts.  The device sampler only sees its result, a per-row byte mask of the tokens to penalise.
#pragma once

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "../../include/whisper.h"

namespace wb {

struct PartialUtf8 {
    uint32_t value = 0;  // bits so far (unshifted)
    int n_remain = 0;    // bytes still missing; -1 = invalid sequence
};

// One decoder's grammar state (whisper_grammar, wc:802-808).  The rule table is immutable and shared between the copies beam search
// makes; a stack is a list of positions inside it, its top always resting on a character-range element.
struct Grammar {
    using Elem = whisper_grammar_element;
    using Stack = std::vector<const Elem *>;
    std::shared_ptr<const std::vector<std::vector<Elem>>> rules;
    std::vector<Stack> stacks;
    PartialUtf8 partial;

    bool active() const { return rules && !rules->empty() && !stacks.empty(); }
    // whisper_grammar_init (wc:5784-5822)
    static Grammar init(const Elem * const * rule_ptrs, size_t n_rules, size_t i_start_rule);
    // whisper_grammar_accept_token (wc:5868-5890): advance over the code points of an accepted token ("[_...]" specials are skipped)
    void accept_token(const std::string & text);
    // whisper_suppress_invalid_grammar (wc:5824-5866): reject[id] = 1 for every token id < eot with non-empty text that no stack
    // can continue with; everything else 0.  `reject` has n_vocab entries.
    void reject_mask(const std::vector<std::string> & id_to_token, int token_eot, std::vector<uint8_t> & reject) const;
};

}  // namespace wb
