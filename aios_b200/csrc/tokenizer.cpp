// tokenizer.cpp — see tokenizer.h.  Host-only; compiled into libaios_b200.so.
#include "tokenizer.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <queue>

#include "gguf.h"

namespace aios {

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

std::string Tokenizer::chat_prompt(const std::string &system, const std::string &user) const {
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
