// tokenizer.cpp — see tokenizer.h.  Host-only; compiled into libaios_b200.so.
#include "tokenizer.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <queue>

#include "gguf.h"
#include "jinja_lite.h"
#include "unicode_tables.h"

namespace aios {

// ---------------------------------------------------------------- Unicode helpers for the BPE pre-tokenizer
namespace {
bool in_ranges(const uint32_t (*tab)[2], int n, uint32_t cp) {
    int lo = 0, hi = n - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) / 2;
        if (cp < tab[mid][0]) hi = mid - 1;
        else if (cp > tab[mid][1]) lo = mid + 1;
        else return true;
    }
    return false;
}
enum CpClass : uint8_t { C_OTHER = 0, C_LETTER = 1, C_NUMBER = 2, C_SPACE = 3 };
CpClass classify(uint32_t cp) {
    if (cp < 128) {
        if ((cp >= 'a' && cp <= 'z') || (cp >= 'A' && cp <= 'Z')) return C_LETTER;
        if (cp >= '0' && cp <= '9') return C_NUMBER;
        if (cp == ' ' || (cp >= 9 && cp <= 13)) return C_SPACE;
        return C_OTHER;
    }
    if (in_ranges(UNI_LETTER, UNI_LETTER_N, cp)) return C_LETTER;
    if (in_ranges(UNI_NUMBER, UNI_NUMBER_N, cp)) return C_NUMBER;
    if (in_ranges(UNI_SPACE, UNI_SPACE_N, cp)) return C_SPACE;
    return C_OTHER;
}

struct Cp {
    uint32_t cp;
    uint32_t off;  // byte offset in the text
    CpClass cls;
};
// decode UTF-8 leniently: an invalid byte is one code point of class OTHER (it still goes through the byte map)
std::vector<Cp> decode_utf8(const std::string &t, size_t off, size_t len) {
    std::vector<Cp> out;
    out.reserve(len);
    const size_t end = off + len;
    for (size_t i = off; i < end;) {
        const unsigned char c = (unsigned char)t[i];
        uint32_t cp = c;
        size_t n = 1;
        if (c >= 0xC2 && c <= 0xDF) n = 2, cp = c & 0x1F;
        else if (c >= 0xE0 && c <= 0xEF) n = 3, cp = c & 0x0F;
        else if (c >= 0xF0 && c <= 0xF4) n = 4, cp = c & 0x07;
        bool ok = n > 1 && i + n <= end;
        for (size_t k = 1; ok && k < n; k++) {
            const unsigned char cc = (unsigned char)t[i + k];
            if ((cc & 0xC0) != 0x80) ok = false;
            else cp = (cp << 6) | (cc & 0x3F);
        }
        if (n > 1 && !ok) {
            out.push_back({0xFFFFFFFFu, (uint32_t)i, C_OTHER});
            i += 1;
            continue;
        }
        out.push_back({cp, (uint32_t)i, n == 1 ? classify(cp) : classify(cp)});
        i += n;
    }
    return out;
}
inline bool is_nl(uint32_t cp) { return cp == '\r' || cp == '\n'; }
inline uint32_t ascii_lower(uint32_t cp) { return cp >= 'A' && cp <= 'Z' ? cp + 32 : cp; }

// Length in code points of the pre-token that starts at i.
//   Llama-3: (?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+
//   GPT-2:   's|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+
// Alternatives are tried in order at each position (leftmost alternative wins, each quantifier greedy with
// backtracking) — written out case by case.
size_t pretoken_len(const std::vector<Cp> &c, size_t i, int pre) {
    const size_t n = c.size();
    const bool l3 = pre == 1;
    // contractions
    if (c[i].cp == '\'' && i + 1 < n) {
        const uint32_t a = l3 ? ascii_lower(c[i + 1].cp) : c[i + 1].cp;
        if (a == 's' || a == 't' || a == 'm' || a == 'd') return 2;
        if (i + 2 < n) {
            const uint32_t b = l3 ? ascii_lower(c[i + 2].cp) : c[i + 2].cp;
            if ((a == 'r' && b == 'e') || (a == 'v' && b == 'e') || (a == 'l' && b == 'l')) return 3;
        }
    }
    auto run = [&](size_t j, CpClass k) {
        while (j < n && c[j].cls == k) j++;
        return j;
    };
    if (l3) {
        // [^\r\n\p{L}\p{N}]?\p{L}+
        if (c[i].cls != C_LETTER && c[i].cls != C_NUMBER && !is_nl(c[i].cp) && i + 1 < n && c[i + 1].cls == C_LETTER)
            return run(i + 1, C_LETTER) - i;
        if (c[i].cls == C_LETTER) return run(i, C_LETTER) - i;
        // \p{N}{1,3}
        if (c[i].cls == C_NUMBER) return std::min(run(i, C_NUMBER), i + 3) - i;
        // ' ?[^\s\p{L}\p{N}]+[\r\n]*'
        {
            size_t j = i;
            if (c[j].cp == ' ' && j + 1 < n && c[j + 1].cls == C_OTHER) j++;
            if (c[j].cls == C_OTHER) {
                j = run(j, C_OTHER);
                while (j < n && is_nl(c[j].cp)) j++;
                return j - i;
            }
        }
    } else {
        // ' ?\p{L}+', ' ?\p{N}+', ' ?[^\s\p{L}\p{N}]+'
        size_t j = i;
        if (c[j].cp == ' ' && j + 1 < n && c[j + 1].cls != C_SPACE) j++;
        if (c[j].cls != C_SPACE) return run(j, c[j].cls) - i;
    }
    // only whitespace can be here
    if (c[i].cls != C_SPACE) return 1;
    const size_t e = run(i, C_SPACE);
    if (l3) {  // \s*[\r\n]+ : up to and including the last newline of the run
        size_t last = n;
        for (size_t j = i; j < e; j++)
            if (is_nl(c[j].cp)) last = j;
        if (last != n) return last + 1 - i;
    }
    // \s+(?!\S): leave the last blank to the word that follows, unless the run is a single one or ends the text
    if (e == n) return e - i;
    if (e - i > 1) return e - 1 - i;
    return e - i;  // \s+
}

// GPT-2's byte -> printable code point map (bytes_to_unicode): printable Latin-1 stays, the rest moves to U+0100...
struct ByteMap {
    std::string enc[256];                          // UTF-8 of the mapped character
    std::unordered_map<uint32_t, uint8_t> dec;     // mapped code point -> byte
    ByteMap() {
        int extra = 0;
        for (int b = 0; b < 256; b++) {
            const bool keep = (b >= 33 && b <= 126) || (b >= 161 && b <= 172) || (b >= 174 && b <= 255);
            const uint32_t cp = keep ? (uint32_t)b : 256u + (uint32_t)extra++;
            if (cp < 0x80) enc[b] = std::string(1, (char)cp);
            else enc[b] = {(char)(0xC0 | (cp >> 6)), (char)(0x80 | (cp & 0x3F))};
            dec.emplace(cp, (uint8_t)b);
        }
    }
};
const ByteMap &byte_map() {
    static const ByteMap m;
    return m;
}
}  // namespace

static const char kSpace[] = "\xE2\x96\x81";  // U+2581, SentencePiece's visible space

static size_t utf8_len(unsigned char c) {
    if (c < 0x80) return 1;
    if ((c & 0xE0) == 0xC0) return 2;
    if ((c & 0xF0) == 0xE0) return 3;
    if ((c & 0xF8) == 0xF0) return 4;
    return 1;  // stray continuation byte: one symbol, ends in byte fallback
}

Tokenizer::Tokenizer(const GgufFile &f) {
    model_ = f.get_s("tokenizer.ggml.model", "");
    pieces_ = f.get_str_array("tokenizer.ggml.tokens");
    types_ = f.get_i32_array("tokenizer.ggml.token_type");
    scores_ = f.get_f32_array("tokenizer.ggml.scores");
    bos_ = (int32_t)f.get_u("tokenizer.ggml.bos_token_id", (uint64_t)-1);
    eos_ = (int32_t)f.get_u("tokenizer.ggml.eos_token_id", (uint64_t)-1);
    unk_ = (int32_t)f.get_u("tokenizer.ggml.unknown_token_id", 0);
    add_bos_ = f.has("tokenizer.ggml.add_bos_token") ? f.get_u("tokenizer.ggml.add_bos_token", 1) != 0 : true;
    add_space_prefix_ = f.has("tokenizer.ggml.add_space_prefix") ? f.get_u("tokenizer.ggml.add_space_prefix", 1) != 0 : true;

    bpe_ = model_ == "gpt2";
    if (bpe_) {
        const std::string pre = f.get_s("tokenizer.ggml.pre", "default");
        if (pre == "llama-bpe" || pre == "llama3" || pre == "llama-v3") {
            pre_ = 1;
            ignore_merges_ = true;  // a pre-token that is itself a piece is taken whole
            if (!f.has("tokenizer.ggml.add_bos_token")) add_bos_ = true;
        } else {
            pre_ = 0;  // GPT-2's pattern (also the fallback for pre-tokenizers this build does not know)
            if (!f.has("tokenizer.ggml.add_bos_token")) add_bos_ = false;
        }
        add_space_prefix_ = false;
        const std::vector<std::string> merges = f.get_str_array("tokenizer.ggml.merges");
        merge_rank_.reserve(merges.size() * 2);
        for (size_t i = 0; i < merges.size(); i++) merge_rank_.emplace(merges[i], (int32_t)i);
    }

    index_.reserve(pieces_.size() * 2);
    for (int i = 0; i < 256; i++) byte_ids_[i] = -1;
    for (int32_t i = 0; i < (int32_t)pieces_.size(); i++) {
        index_.emplace(pieces_[i], i);  // first id wins for duplicate strings
        int t = type(i);
        if (t == TOK_BYTE && pieces_[i].size() == 6 && pieces_[i].compare(0, 3, "<0x") == 0 && pieces_[i][5] == '>') {
            unsigned v = 0;
            if (sscanf(pieces_[i].c_str() + 3, "%2X", &v) == 1) byte_ids_[v & 255] = i;
        }
        if ((t == TOK_CONTROL || t == TOK_USER_DEFINED) && !pieces_[i].empty()) specials_.push_back(i);
    }
    std::stable_sort(specials_.begin(), specials_.end(),
                     [&](int32_t a, int32_t b) { return pieces_[a].size() > pieces_[b].size(); });

    if (eos_ >= 0) eog_.push_back(eos_);
    for (const char *s : {"<|im_end|>", "<|eot_id|>", "<|end_of_text|>", "<|endoftext|>", "<|end|>", "<end_of_turn>"}) {
        auto it = index_.find(s);
        if (it != index_.end() && std::find(eog_.begin(), eog_.end(), it->second) == eog_.end()) eog_.push_back(it->second);
    }

    // chat format from the marker strings of tokenizer.chat_template; chatml when there is none
    const std::string t = f.get_s("tokenizer.chat_template", "");
    template_ = t;
    auto has = [&](const char *s) { return t.find(s) != std::string::npos; };
    if (has("<|im_start|>")) format_ = CHAT_CHATML;
    else if (has("<|start_header_id|>")) format_ = CHAT_LLAMA3;
    else if (has("[INST]")) format_ = has("<<SYS>>") ? CHAT_LLAMA2 : CHAT_MISTRAL;
    else if (has("<|user|>")) format_ = CHAT_ZEPHYR;
    else format_ = CHAT_CHATML;
}

bool Tokenizer::is_eog(int32_t id) const { return std::find(eog_.begin(), eog_.end(), id) != eog_.end(); }

// one final symbol -> its id, or one byte token per byte when the vocabulary has no such piece
void Tokenizer::emit_symbol(const std::string &s, std::vector<int32_t> &out) const {
    auto it = index_.find(s);
    if (it != index_.end()) {
        out.push_back(it->second);
        return;
    }
    for (unsigned char c : s) out.push_back(byte_ids_[c] >= 0 ? byte_ids_[c] : unk_);
}

namespace {
struct Symbol {
    int prev, next;
    size_t off, len;  // slice of the text; len 0 = merged away
};
struct Bigram {
    int left, right;
    float score;
    size_t size;
    bool operator<(const Bigram &o) const {  // max-heap on score, leftmost first on ties
        return score < o.score || (score == o.score && left > o.left);
    }
};
}  // namespace

void Tokenizer::encode_spm(const std::string &text, std::vector<int32_t> &out) const {
    if (text.empty()) return;
    std::vector<Symbol> sym;
    for (size_t off = 0; off < text.size();) {
        size_t n = std::min(utf8_len((unsigned char)text[off]), text.size() - off);
        sym.push_back({(int)sym.size() - 1, (int)sym.size() + 1, off, n});
        off += n;
    }
    sym.back().next = -1;

    // a pair is a merge candidate only when its concatenation is itself a piece, so every surviving
    // symbol is either one character or a vocabulary piece
    std::priority_queue<Bigram> work;
    auto try_add = [&](int l, int r) {
        if (l < 0 || r < 0) return;
        std::string s = text.substr(sym[l].off, sym[l].len + sym[r].len);
        auto it = index_.find(s);
        if (it == index_.end()) return;
        work.push({l, r, score(it->second), s.size()});
    };
    for (int i = 1; i < (int)sym.size(); i++) try_add(i - 1, i);

    while (!work.empty()) {
        Bigram b = work.top();
        work.pop();
        Symbol &l = sym[b.left], &r = sym[b.right];
        if (l.len == 0 || r.len == 0 || l.len + r.len != b.size) continue;  // stale entry
        l.len += r.len;
        r.len = 0;
        l.next = r.next;
        if (r.next >= 0) sym[r.next].prev = b.left;
        try_add(l.prev, b.left);
        try_add(b.left, l.next);
    }
    for (int i = 0; i >= 0; i = sym[i].next) emit_symbol(text.substr(sym[i].off, sym[i].len), out);
}

// one pre-token, already in mapped characters: merge adjacent symbols by rank until no listed pair is left
void Tokenizer::bpe_word(const std::string &w, std::vector<int32_t> &out) const {
    if (ignore_merges_) {
        auto it = index_.find(w);
        if (it != index_.end()) {
            out.push_back(it->second);
            return;
        }
    }
    std::vector<Symbol> sym;
    for (size_t off = 0; off < w.size();) {
        const size_t n = std::min(utf8_len((unsigned char)w[off]), w.size() - off);
        sym.push_back({(int)sym.size() - 1, (int)sym.size() + 1, off, n});
        off += n;
    }
    sym.back().next = -1;
    struct Pair {
        int left, right, rank;
        size_t llen, rlen;
        bool operator<(const Pair &o) const { return rank > o.rank || (rank == o.rank && left > o.left); }  // min-heap
    };
    std::priority_queue<Pair> work;
    std::string key;
    auto try_add = [&](int l, int r) {
        if (l < 0 || r < 0) return;
        key.assign(w, sym[l].off, sym[l].len);
        key += ' ';
        key.append(w, sym[r].off, sym[r].len);
        auto it = merge_rank_.find(key);
        if (it != merge_rank_.end()) work.push({l, r, it->second, sym[l].len, sym[r].len});
    };
    for (int i = 1; i < (int)sym.size(); i++) try_add(i - 1, i);
    while (!work.empty()) {
        const Pair b = work.top();
        work.pop();
        Symbol &l = sym[b.left], &r = sym[b.right];
        if (l.len != b.llen || r.len != b.rlen || l.next != b.right) continue;  // stale entry
        l.len += r.len;
        r.len = 0;
        l.next = r.next;
        if (r.next >= 0) sym[r.next].prev = b.left;
        try_add(l.prev, b.left);
        try_add(b.left, l.next);
    }
    for (int i = 0; i >= 0; i = sym[i].next) {
        const std::string piece = w.substr(sym[i].off, sym[i].len);
        auto it = index_.find(piece);
        if (it != index_.end()) {
            out.push_back(it->second);
            continue;
        }
        for (size_t off = 0; off < piece.size();) {  // no such piece: one token per byte (mapped character)
            const size_t n = std::min(utf8_len((unsigned char)piece[off]), piece.size() - off);
            auto b = index_.find(piece.substr(off, n));
            out.push_back(b != index_.end() ? b->second : unk_);
            off += n;
        }
    }
}

void Tokenizer::encode_bpe(const std::string &text, std::vector<int32_t> &out) const {
    const std::vector<Cp> cps = decode_utf8(text, 0, text.size());
    const ByteMap &bm = byte_map();
    std::string w;
    for (size_t i = 0; i < cps.size();) {
        const size_t len = pretoken_len(cps, i, pre_);
        const size_t b0 = cps[i].off, b1 = i + len < cps.size() ? cps[i + len].off : text.size();
        w.clear();
        for (size_t b = b0; b < b1; b++) w += bm.enc[(unsigned char)text[b]];
        bpe_word(w, out);
        i += len;
    }
}

std::vector<int32_t> Tokenizer::encode(const std::string &text, bool add_special, bool parse_special) const {
    std::vector<int32_t> out;
    if (add_special && add_bos_ && bos_ >= 0) out.push_back(bos_);

    // split the text at special-token strings (longest first), keep raw fragments in between
    struct Frag { bool special; int32_t id; size_t off, len; };
    std::vector<Frag> frags{{false, -1, 0, text.size()}};
    if (parse_special) {
        for (int32_t id : specials_) {
            const std::string &p = pieces_[id];
            std::vector<Frag> next;
            for (const Frag &fr : frags) {
                if (fr.special) { next.push_back(fr); continue; }
                size_t cur = fr.off, end = fr.off + fr.len;
                while (cur < end) {
                    size_t hit = text.find(p, cur);
                    if (hit == std::string::npos || hit + p.size() > end) break;
                    if (hit > cur) next.push_back({false, -1, cur, hit - cur});
                    next.push_back({true, id, hit, p.size()});
                    cur = hit + p.size();
                }
                if (cur < end) next.push_back({false, -1, cur, end - cur});
            }
            frags.swap(next);
        }
    }

    bool prev_special = true;  // the start of the text counts as a boundary
    for (const Frag &fr : frags) {
        if (fr.special) {
            out.push_back(fr.id);
            prev_special = true;
            continue;
        }
        if (fr.len == 0) continue;
        if (bpe_) {
            encode_bpe(text.substr(fr.off, fr.len), out);
            prev_special = false;
            continue;
        }
        std::string s;
        s.reserve(fr.len * 3 + 3);
        if (add_space_prefix_ && prev_special) s += kSpace;
        for (size_t i = fr.off; i < fr.off + fr.len; i++) {
            if (text[i] == ' ') s += kSpace;
            else s += text[i];
        }
        encode_spm(s, out);
        prev_special = false;
    }
    return out;
}

std::string Tokenizer::piece_text(int32_t id, bool render_special) const {
    if (id < 0 || id >= n_tokens()) return std::string();
    const std::string &p = pieces_[id];
    switch (type(id)) {
        case TOK_BYTE: {
            unsigned v = 0;
            if (p.size() == 6 && sscanf(p.c_str() + 3, "%2X", &v) == 1) return std::string(1, (char)v);
            return std::string();
        }
        case TOK_CONTROL:
            return render_special ? p : std::string();
        case TOK_UNKNOWN:
            return render_special ? p : std::string();
        case TOK_UNUSED:
            return std::string();
        default: {
            std::string s;
            s.reserve(p.size());
            if (bpe_) {  // mapped characters back to bytes (user-defined pieces are literal text)
                if (type(id) == TOK_USER_DEFINED) return p;
                const ByteMap &bm = byte_map();
                for (const Cp &c : decode_utf8(p, 0, p.size())) {
                    auto it = bm.dec.find(c.cp);
                    if (it != bm.dec.end()) s += (char)it->second;
                }
                return s;
            }
            for (size_t i = 0; i < p.size();) {
                if (p.compare(i, 3, kSpace) == 0) { s += ' '; i += 3; }
                else s += p[i++];
            }
            return s;
        }
    }
}

std::string Tokenizer::decode(const int32_t *ids, int32_t n, bool render_special, bool strip_space) const {
    std::string s;
    for (int32_t i = 0; i < n; i++) s += piece_text(ids[i], render_special);
    if (strip_space && add_space_prefix_ && !s.empty() && s[0] == ' ') s.erase(0, 1);
    return s;
}

bool Tokenizer::apply_template(const std::vector<std::pair<std::string, std::string>> &messages, bool add_generation_prompt,
                               std::string *out, std::string *err) const {
    if (template_.empty()) {
        if (err) *err = "the file has no tokenizer.chat_template";
        return false;
    }
    std::vector<ChatMessage> msgs;
    for (const auto &m : messages) msgs.push_back({m.first, m.second});
    const std::string bos_text = bos_ >= 0 && bos_ < n_tokens() ? pieces_[bos_] : std::string();
    const std::string eos_text = eos_ >= 0 && eos_ < n_tokens() ? pieces_[eos_] : std::string();
    if (!render_chat_template(template_, msgs, bos_text, eos_text, add_generation_prompt, out, err)) return false;
    // the tokenizer prepends BOS itself (add_special): a BOS string the template emits first would make it two
    if (add_bos_ && !bos_text.empty() && out->compare(0, bos_text.size(), bos_text) == 0) out->erase(0, bos_text.size());
    return true;
}

std::string Tokenizer::chat_prompt(const std::string &system, const std::string &user) const {
    {
        std::vector<std::pair<std::string, std::string>> msgs;
        if (!system.empty()) msgs.emplace_back("system", system);
        msgs.emplace_back("user", user);
        std::string rendered, err;
        if (apply_template(msgs, true, &rendered, &err)) return rendered;
    }
    std::string s;
    const std::string eos_text = eos_ >= 0 && eos_ < n_tokens() ? pieces_[eos_] : std::string("</s>");
    switch (format_) {
        case CHAT_LLAMA2:
            s = "[INST] ";
            if (!system.empty()) s += "<<SYS>>\n" + system + "\n<</SYS>>\n\n";
            s += user + " [/INST]";
            break;
        case CHAT_MISTRAL:
            s = "[INST] ";
            if (!system.empty()) s += system + "\n\n";
            s += user + " [/INST]";
            break;
        case CHAT_ZEPHYR:
            if (!system.empty()) s += "<|system|>\n" + system + eos_text + "\n";
            s += "<|user|>\n" + user + eos_text + "\n<|assistant|>\n";
            break;
        case CHAT_LLAMA3:
            if (!system.empty()) s += "<|start_header_id|>system<|end_header_id|>\n\n" + system + "<|eot_id|>";
            s += "<|start_header_id|>user<|end_header_id|>\n\n" + user + "<|eot_id|>";
            s += "<|start_header_id|>assistant<|end_header_id|>\n\n";
            break;
        case CHAT_CHATML:
        default:
            if (!system.empty()) s += "<|im_start|>system\n" + system + "<|im_end|>\n";
            s += "<|im_start|>user\n" + user + "<|im_end|>\n<|im_start|>assistant\n";
            break;
    }
    return s;
}

}  // namespace aios
