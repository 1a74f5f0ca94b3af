// json_mask.cpp — JSON mode on the device side of the sampler.
//
// The reference's unary Infer always asks llama-server for `response_format: json_object`
// (runtime/src/inference.rs:119-121) and the orchestrator parses the answer as JSON
// (agent-core/src/autonomy.rs:839-845).  llama-server implements that as a grammar (`root ::= object`) that removes
// candidates before its sampler chain.  Here the grammar is a byte-level pushdown automaton (the same one as
// aios_b200/runtime/json_grammar.py, restated in C++ so that no Python runs per token), walked over every token's bytes
// once per step to produce the allowed set as a bit mask; the mask goes to the GPU (n_vocab / 8 bytes) and
// sample_kernel applies it.  aios_json_grammar_mask_cb has the aios_mask_cb signature: hand it, with the grammar as
// `user`, to aios_request_submit_masked and the scheduler thread drives it between decode steps.
//
// Grammar (RFC 8259 as llama.cpp's json.gbnf restricts it): root = ws object ws; value = object | array | string |
// number | true | false | null; strings with the standard escapes and \uXXXX; numbers -?(0|[1-9][0-9]*)(.[0-9]+)?
// ([eE][+-]?[0-9]+)?; whitespace = space, tab, CR, LF; bytes >= 0x80 inside strings must form well-formed UTF-8.
#include "json_mask.h"

#include <string.h>

namespace aios {

namespace {
enum Mode : uint8_t { M_VALUE, M_OBJ_FIRST_KEY, M_OBJ_KEY, M_OBJ_COLON, M_OBJ_NEXT, M_ARR_FIRST, M_ARR_NEXT, M_STRING, M_STR_ESC,
                      M_STR_HEX, M_NUMBER, M_LITERAL, M_DONE, M_START };
enum Num : uint8_t { N_MINUS, N_ZERO, N_INT, N_DOT, N_FRAC, N_E, N_ESIGN, N_EXP };
const char *const LITS[3] = {"true", "false", "null"};
inline bool is_ws(uint8_t b) { return b == ' ' || b == '\t' || b == '\r' || b == '\n'; }
inline bool is_digit(uint8_t b) { return b >= '0' && b <= '9'; }
inline bool is_hex(uint8_t b) { return is_digit(b) || (b >= 'a' && b <= 'f') || (b >= 'A' && b <= 'F'); }
}  // namespace

void JsonPrefix::after_value() {
    if (depth == 0) mode = M_DONE;
    else mode = in_array(depth - 1) ? M_ARR_NEXT : M_OBJ_NEXT;
}

bool JsonPrefix::push(bool array) {
    if (depth >= MAX_DEPTH) return false;
    if (array) stack[depth >> 6] |= (uint64_t)1 << (depth & 63);
    else stack[depth >> 6] &= ~((uint64_t)1 << (depth & 63));
    depth++;
    return true;
}

bool JsonPrefix::start_value(uint8_t b) {
    switch (b) {
        case '{': if (!push(false)) return false; mode = M_OBJ_FIRST_KEY; return true;
        case '[': if (!push(true)) return false; mode = M_ARR_FIRST; return true;
        case '"': mode = M_STRING; key_string = false; return true;
        case '-': mode = M_NUMBER; sub = N_MINUS; return true;
        case '0': mode = M_NUMBER; sub = N_ZERO; return true;
        case 't': mode = M_LITERAL; lit = 0; sub = 1; return true;
        case 'f': mode = M_LITERAL; lit = 1; sub = 1; return true;
        case 'n': mode = M_LITERAL; lit = 2; sub = 1; return true;
        default: break;
    }
    if (b >= '1' && b <= '9') { mode = M_NUMBER; sub = N_INT; return true; }
    return false;
}

bool JsonPrefix::byte(uint8_t b) {
    switch (mode) {
        case M_STRING:
            if (utf8_left) {
                if (b >= 0x80 && b <= 0xBF) { utf8_left--; return true; }
                return false;
            }
            if (b == '"') {
                if (key_string) mode = M_OBJ_COLON;
                else after_value();
                return true;
            }
            if (b == '\\') { mode = M_STR_ESC; return true; }
            if (b < 0x20) return false;
            if (b < 0x80) return true;
            if (b >= 0xC2 && b <= 0xDF) utf8_left = 1;
            else if (b >= 0xE0 && b <= 0xEF) utf8_left = 2;
            else if (b >= 0xF0 && b <= 0xF4) utf8_left = 3;
            else return false;
            return true;
        case M_STR_ESC:
            if (b == '"' || b == '\\' || b == '/' || b == 'b' || b == 'f' || b == 'n' || b == 'r' || b == 't') { mode = M_STRING; return true; }
            if (b == 'u') { mode = M_STR_HEX; sub = 4; return true; }
            return false;
        case M_STR_HEX:
            if (!is_hex(b)) return false;
            if (--sub == 0) mode = M_STRING;
            return true;
        case M_NUMBER: {
            const uint8_t s = sub;
            if (s == N_MINUS) {
                if (b == '0') { sub = N_ZERO; return true; }
                if (b >= '1' && b <= '9') { sub = N_INT; return true; }
                return false;
            }
            if ((s == N_INT || s == N_FRAC || s == N_EXP) && is_digit(b)) return true;
            if ((s == N_ZERO || s == N_INT) && b == '.') { sub = N_DOT; return true; }
            if ((s == N_ZERO || s == N_INT || s == N_FRAC) && (b == 'e' || b == 'E')) { sub = N_E; return true; }
            if (s == N_DOT) {
                if (is_digit(b)) { sub = N_FRAC; return true; }
                return false;
            }
            if (s == N_E) {
                if (b == '+' || b == '-') { sub = N_ESIGN; return true; }
                if (is_digit(b)) { sub = N_EXP; return true; }
                return false;
            }
            if (s == N_ESIGN) {
                if (is_digit(b)) { sub = N_EXP; return true; }
                return false;
            }
            // anything else ends the number, if it may end here, and is then looked at again
            if (!(s == N_ZERO || s == N_INT || s == N_FRAC || s == N_EXP)) return false;
            after_value();
            return byte(b);
        }
        case M_LITERAL:
            if (b != (uint8_t)LITS[lit][sub]) return false;
            if (LITS[lit][++sub] == 0) after_value();
            return true;
        default: break;
    }
    // structural modes: whitespace is free
    if (is_ws(b)) return true;
    switch (mode) {
        case M_START:
            if (b == '{') { push(false); mode = M_OBJ_FIRST_KEY; return true; }
            return false;
        case M_VALUE: return start_value(b);
        case M_OBJ_FIRST_KEY:
        case M_OBJ_KEY:
            if (b == '"') { mode = M_STRING; key_string = true; return true; }
            if (b == '}' && mode == M_OBJ_FIRST_KEY) { depth--; after_value(); return true; }
            return false;
        case M_OBJ_COLON:
            if (b == ':') { mode = M_VALUE; return true; }
            return false;
        case M_OBJ_NEXT:
            if (b == ',') { mode = M_OBJ_KEY; return true; }
            if (b == '}') { depth--; after_value(); return true; }
            return false;
        case M_ARR_FIRST:
            if (b == ']') { depth--; after_value(); return true; }
            return start_value(b);
        case M_ARR_NEXT:
            if (b == ',') { mode = M_VALUE; return true; }
            if (b == ']') { depth--; after_value(); return true; }
            return false;
        default: return false;  // M_DONE: only whitespace (handled above)
    }
}

JsonPrefix::JsonPrefix() { memset(this, 0, sizeof(*this)); mode = M_START; }
bool JsonPrefix::complete() const { return mode == M_DONE; }

bool JsonPrefix::feed(const uint8_t *data, size_t n) {
    for (size_t i = 0; i < n; i++)
        if (!byte(data[i])) return false;
    return true;
}

uint64_t JsonPrefix::key(int levels) const {
    if (levels > 12 || depth > MAX_DEPTH - 20) return 0;  // deep pops / near the depth limit: compute afresh
    // a piece with `levels` closing brackets pops that many containers and then continues inside the one below: the
    // key holds the kinds of the top levels + 1 containers, the container count up to levels + 2, and the scalar state
    uint64_t top = 0;
    const int n = depth < levels + 1 ? depth : levels + 1;
    for (int i = 0; i < n; i++) top |= (uint64_t)in_array(depth - 1 - i) << i;
    const uint64_t dcap = (uint64_t)(depth < levels + 2 ? depth : levels + 2);
    return 1 | (uint64_t)mode << 1 | (uint64_t)sub << 6 | (uint64_t)lit << 10 | (uint64_t)utf8_left << 12 | (uint64_t)key_string << 14 |
           dcap << 15 | top << 20;
}

JsonTables::JsonTables(const Tokenizer &tok) {
    const int32_t n = tok.n_tokens();
    pieces.resize((size_t)n);
    for (int32_t i = 0; i < n; i++) {
        pieces[(size_t)i] = tok.decode(&i, 1, false, false);  // control tokens: empty -> never allowed
        int c = 0;
        for (char ch : pieces[(size_t)i]) c += ch == '}' || ch == ']';
        if (c > max_closers) max_closers = c;
    }
}

int JsonMasker::step(int32_t last_token, uint8_t *mask, int32_t n_bytes) {
    const std::vector<std::string> &pieces = tab->pieces;
    if (last_token >= 0) {
        if ((size_t)last_token >= pieces.size()) return -1;
        const std::string &p = pieces[(size_t)last_token];
        if (p.empty() || !state.feed(reinterpret_cast<const uint8_t *>(p.data()), p.size())) return -1;  // cannot happen under the mask
    }
    if (state.complete()) return 2;  // the object has closed
    const int32_t n = (int32_t)pieces.size();
    const int32_t need = (n + 7) / 8;
    const uint64_t k = n_bytes >= need ? state.key(tab->max_closers) : 0;
    if (k) {
        std::lock_guard<std::mutex> lk(tab->mu);
        auto it = tab->cache.find(k);
        if (it != tab->cache.end()) {
            memcpy(mask, it->second.data(), (size_t)need);
            return it->second[(size_t)need] ? 0 : 2;
        }
    }
    int n_allowed = 0;
    for (int32_t i = 0; i < n && (i >> 3) < n_bytes; i++) {
        const std::string &p = pieces[(size_t)i];
        if (p.empty()) continue;
        JsonPrefix s = state;  // plain struct copy: try the piece on a scratch state
        if (s.feed(reinterpret_cast<const uint8_t *>(p.data()), p.size())) {
            mask[i >> 3] |= (uint8_t)(1u << (i & 7));
            n_allowed++;
        }
    }
    if (k) {
        std::vector<uint8_t> v(mask, mask + need);
        v.push_back(n_allowed ? 1 : 0);
        std::lock_guard<std::mutex> lk(tab->mu);
        if (tab->cache.size() < 4096) tab->cache.emplace(k, std::move(v));
    }
    return n_allowed ? 0 : 2;  // nothing fits: end the request rather than sample outside the grammar
}

}  // namespace aios
