// Grammar-constrained sampling: the pushdown automaton behind whisper_full_params.grammar_rules (reference: whisper.cpp:5484-5896).
// Same observable behaviour as the reference (which tokens are penalised at every step), own structure: positions are pointers into an
// immutable shared rule table, the candidate walk is iterative per stack, and the result is a byte mask for the device sampler.
#include "grammar.h"

#include <string.h>

namespace wb {

namespace {

using Elem = Grammar::Elem;
using Stack = Grammar::Stack;
using Rules = std::vector<std::vector<Elem>>;

inline bool ends_alternative(const Elem * e) { return e->type == WHISPER_GRETYPE_END || e->type == WHISPER_GRETYPE_ALT; }

// code points of `text` continuing a possibly incomplete sequence; a 0 terminates the list (wc:5484-5538)
struct Decoded {
    std::vector<uint32_t> cps;
    PartialUtf8 tail;
};
Decoded decode_utf8(const char * s, PartialUtf8 start) {
    static const int n_bytes_by_high_nibble[16] = {1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 2, 2, 3, 4};
    Decoded out;
    uint32_t value = start.value;
    int remain = start.n_remain;
    for (; *s != 0 && remain > 0; ++s, --remain) {          // finish the sequence the previous token left open
        const uint8_t b = (uint8_t) *s;
        if ((b >> 6) != 2) {
            out.cps.push_back(0);
            out.tail = PartialUtf8{0, -1};
            return out;
        }
        value = (value << 6) + (b & 0x3F);
    }
    if (start.n_remain > 0 && remain == 0) out.cps.push_back(value);
    while (*s != 0) {
        const uint8_t first = (uint8_t) *s;
        remain = n_bytes_by_high_nibble[first >> 4] - 1;
        if (remain < 0) {                                    // a continuation byte where a sequence must start
            out.cps.assign(1, 0);
            out.tail = PartialUtf8{0, remain};
            return out;
        }
        value = first & ((1u << (7 - remain)) - 1);
        ++s;
        for (; *s != 0 && remain > 0; ++s, --remain) value = (value << 6) + ((uint8_t) *s & 0x3F);
        if (remain == 0) out.cps.push_back(value);
    }
    out.cps.push_back(0);
    out.tail = PartialUtf8{value, remain};
    return out;
}

// does `chr` satisfy the character class starting at `pos` (CHAR or CHAR_NOT, then CHAR_ALT / CHAR_RNG_UPPER continuations)?
// *after receives the element following the class (wc:5551-5573)
bool class_matches(const Elem * pos, uint32_t chr, const Elem ** after) {
    const bool positive = pos->type == WHISPER_GRETYPE_CHAR;
    bool hit = false;
    do {
        if (pos[1].type == WHISPER_GRETYPE_CHAR_RNG_UPPER) {
            hit = hit || (pos->value <= chr && chr <= pos[1].value);
            pos += 2;
        } else {
            hit = hit || pos->value == chr;
            pos += 1;
        }
    } while (pos->type == WHISPER_GRETYPE_CHAR_ALT);
    if (after) *after = pos;
    return hit == positive;
}

// could SOME completion of an unfinished UTF-8 sequence satisfy the class at `pos`? (wc:5578-5622)
bool class_may_match_partial(const Elem * pos, PartialUtf8 p) {
    const bool positive = pos->type == WHISPER_GRETYPE_CHAR;
    if (p.n_remain < 0 || (p.n_remain == 1 && p.value < 2)) return false;  // invalid, or an overlong 2-byte form
    uint32_t low = p.value << (p.n_remain * 6);
    const uint32_t high = low | ((1u << (p.n_remain * 6)) - 1);
    if (low == 0) {
        if (p.n_remain == 2) low = 1u << 11;
        else if (p.n_remain == 3) low = 1u << 16;
    }
    do {
        if (pos[1].type == WHISPER_GRETYPE_CHAR_RNG_UPPER) {
            if (pos->value <= high && low <= pos[1].value) return positive;
            pos += 2;
        } else {
            if (low <= pos->value && pos->value <= high) return positive;
            pos += 1;
        }
    } while (pos->type == WHISPER_GRETYPE_CHAR_ALT);
    return !positive;
}

// expand rule references on top of `stack` until every resulting stack rests on a character class (or is empty) (wc:5627-5678)
void expand(const Rules & rules, const Stack & stack, std::vector<Stack> & out) {
    if (stack.empty()) {
        out.emplace_back();
        return;
    }
    const Elem * top = stack.back();
    if (top->type == WHISPER_GRETYPE_CHAR || top->type == WHISPER_GRETYPE_CHAR_NOT) {
        out.push_back(stack);
        return;
    }
    if (top->type != WHISPER_GRETYPE_RULE_REF) return;  // malformed table: a stack never rests on END / ALT / range continuations
    const Elem * alt = rules[(size_t) top->value].data();
    while (true) {
        Stack next(stack.begin(), stack.end() - 1);
        if (!ends_alternative(top + 1)) next.push_back(top + 1);   // what follows the reference, to come back to
        if (!ends_alternative(alt)) next.push_back(alt);           // this alternative's first element
        expand(rules, next, out);
        while (!ends_alternative(alt)) ++alt;
        if (alt->type != WHISPER_GRETYPE_ALT) break;
        ++alt;
    }
}

// the stack after its top character class has consumed one character
Stack pop_class(const Stack & stack, const Elem * after) {
    Stack next(stack.begin(), stack.end() - 1);
    if (!ends_alternative(after)) next.push_back(after);
    return next;
}

struct Candidate {
    int id;
    const uint32_t * cp;   // remaining code points, 0-terminated
    PartialUtf8 tail;
};

std::vector<Candidate> rejected_by_all(const Rules & rules, const std::vector<Stack> & stacks, const std::vector<Candidate> & cands);

// candidates that `stack` cannot continue with (wc:5717-5766)
std::vector<Candidate> rejected_by(const Rules & rules, const Stack & stack, const std::vector<Candidate> & cands) {
    std::vector<Candidate> rejects;
    if (stack.empty()) {  // the grammar is complete here: only an empty remainder is acceptable
        for (const Candidate & c : cands)
            if (*c.cp != 0 || c.tail.n_remain != 0) rejects.push_back(c);
        return rejects;
    }
    const Elem * top = stack.back();
    std::vector<Candidate> advanced;
    for (const Candidate & c : cands) {
        if (*c.cp == 0) {  // all complete code points consumed: an unfinished tail must still be able to satisfy this class
            if (c.tail.n_remain != 0 && !class_may_match_partial(top, c.tail)) rejects.push_back(c);
        } else if (class_matches(top, *c.cp, nullptr)) {
            advanced.push_back(Candidate{c.id, c.cp + 1, c.tail});
        } else {
            rejects.push_back(c);
        }
    }
    const Elem * after = nullptr;
    class_matches(top, 0, &after);
    std::vector<Stack> next_stacks;
    expand(rules, pop_class(stack, after), next_stacks);
    for (const Candidate & c : rejected_by_all(rules, next_stacks, advanced)) rejects.push_back(Candidate{c.id, c.cp - 1, c.tail});
    return rejects;
}

// a candidate is rejected only if EVERY stack rejects it (wc:5768-5782)
std::vector<Candidate> rejected_by_all(const Rules & rules, const std::vector<Stack> & stacks, const std::vector<Candidate> & cands) {
    if (cands.empty() || stacks.empty()) return {};
    std::vector<Candidate> rejects = rejected_by(rules, stacks.front(), cands);
    for (size_t i = 1; i < stacks.size(); ++i) rejects = rejected_by(rules, stacks[i], rejects);
    return rejects;
}

}  // namespace

Grammar Grammar::init(const Elem * const * rule_ptrs, size_t n_rules, size_t i_start_rule) {
    Grammar g;
    if (!rule_ptrs || n_rules == 0 || i_start_rule >= n_rules) return g;
    auto table = std::make_shared<Rules>(n_rules);
    for (size_t i = 0; i < n_rules; ++i) {
        for (const Elem * e = rule_ptrs[i]; e->type != WHISPER_GRETYPE_END; ++e) (*table)[i].push_back(*e);
        (*table)[i].push_back(Elem{WHISPER_GRETYPE_END, 0});
    }
    g.rules = table;
    const Elem * alt = (*table)[i_start_rule].data();   // one initial stack per alternative of the start rule
    while (true) {
        Stack s;
        if (!ends_alternative(alt)) s.push_back(alt);
        expand(*table, s, g.stacks);
        while (!ends_alternative(alt)) ++alt;
        if (alt->type != WHISPER_GRETYPE_ALT) break;
        ++alt;
    }
    return g;
}

void Grammar::accept_token(const std::string & text) {
    if (!active()) return;
    if (text.rfind("[_", 0) == 0) return;  // special tokens do not move the grammar
    const Decoded d = decode_utf8(text.c_str(), partial);
    for (size_t k = 0; k + 1 < d.cps.size(); ++k) {   // the last entry is the terminator
        std::vector<Stack> next;
        for (const Stack & s : stacks) {
            if (s.empty()) continue;
            const Elem * after = nullptr;
            if (class_matches(s.back(), d.cps[k], &after)) expand(*rules, pop_class(s, after), next);
        }
        stacks.swap(next);
    }
    partial = d.tail;
}

void Grammar::reject_mask(const std::vector<std::string> & id_to_token, int token_eot, std::vector<uint8_t> & reject) const {
    reject.assign(id_to_token.size(), 0);
    if (!active()) return;
    std::vector<Decoded> decoded;
    decoded.reserve((size_t) token_eot);
    std::vector<int> ids;
    for (int id = 0; id < token_eot && id < (int) id_to_token.size(); ++id) {
        if (id_to_token[(size_t) id].empty()) continue;
        decoded.push_back(decode_utf8(id_to_token[(size_t) id].c_str(), partial));
        ids.push_back(id);
    }
    std::vector<Candidate> cands;
    cands.reserve(ids.size());
    for (size_t k = 0; k < ids.size(); ++k) cands.push_back(Candidate{ids[k], decoded[k].cps.data(), decoded[k].tail});
    for (const Candidate & c : rejected_by_all(*rules, stacks, cands)) reject[(size_t) c.id] = 1;
}

}  // namespace wb
