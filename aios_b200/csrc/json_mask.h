// json_mask.h — JSON-object grammar as a byte-level pushdown automaton + the per-step allowed-token mask (json_mask.cpp).
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "tokenizer.h"

namespace aios {

// Is the text so far a prefix of ONE JSON object?  A plain struct (copy = clone), so candidates are tried on a copy.
struct JsonPrefix {
    static constexpr int MAX_DEPTH = 128;
    uint64_t stack[MAX_DEPTH / 64];  // bit d: container d is an array (else an object)
    int32_t depth;
    uint8_t mode, sub, lit, utf8_left;
    bool key_string;

    JsonPrefix();
    bool feed(const uint8_t *data, size_t n);  // false: cannot extend a JSON object (state then undefined)
    bool complete() const;                      // the top-level object has closed

    // cache key of the state as far as a piece with at most `levels` closing brackets can see it; 0 = do not cache
    uint64_t key(int levels) const;

   private:
    bool byte(uint8_t b);
    bool start_value(uint8_t b);
    bool push(bool array);
    void after_value();
    bool in_array(int d) const { return (stack[d >> 6] >> (d & 63)) & 1; }
};

// Per-vocabulary tables shared by every request: the detokenised bytes of every token and a cache of finished masks.
// A mask depends on the automaton state only through (mode, sub-state, the top few containers, whether more containers
// lie below), so the states a generation spends most of its time in — inside a string, after a comma — are computed
// once per vocabulary and then cost one 4 KB copy per step.
struct JsonTables {
    std::vector<std::string> pieces;  // "" for control tokens: never allowed
    int max_closers = 0;              // most '}' / ']' bytes in one piece: how deep a piece can pop the stack
    std::mutex mu;
    std::unordered_map<uint64_t, std::vector<uint8_t>> cache;  // state key -> (mask bytes, then one byte: any token allowed)
    explicit JsonTables(const Tokenizer &tok);
};

// One request's grammar state over one vocabulary: step() advances by the token sampled last and writes the allowed
// set of the next token.
struct JsonMasker {
    std::shared_ptr<JsonTables> tab;
    JsonPrefix state;
    explicit JsonMasker(std::shared_ptr<JsonTables> t) : tab(std::move(t)) {}
    // mask: n_bytes zeroed bytes, bit i of byte i/8 = token i allowed.  Returns 0 (mask written), 2 (object closed or
    // nothing fits: finish the request), -1 (last_token is not a token the grammar allowed).
    int step(int32_t last_token, uint8_t *mask, int32_t n_bytes);
};

}  // namespace aios
