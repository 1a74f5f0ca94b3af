// tokenizer.h — text <-> token ids from the GGUF's tokenizer.ggml.* metadata (host side, C++17).
//
// Replaces the tokenizer / chat-template stage llama-server performs for the reference
// (SURVEY.md §8a D2, §8f-2): the reference sends {system,user} messages
// (runtime/src/inference.rs:363-376) and llama-server renders them with the model's chat
// template and encodes the text.  llama.cpp is not vendored, so this is written from the
// published algorithms [upstream-knowledge]: SentencePiece-style greedy bigram merging by
// score with <0xNN> byte fallback ("llama" vocabularies: TinyLlama, Mistral); byte-level BPE
// ("gpt2" vocabularies: Llama-3) — regex pre-tokenizer (the Llama-3 and GPT-2 patterns, matched by
// a hand-written scanner over Unicode categories), GPT-2's byte <-> printable-character map, merges
// by rank from tokenizer.ggml.merges; special-token partitioning; and recognition of the common
// chat formats by their marker strings (how llama.cpp's llama_chat_apply_template picks one too).
#pragma once
#include <stdint.h>

#include <string>
#include <unordered_map>
#include <vector>

namespace aios {

class GgufFile;

enum TokenType { TOK_NORMAL = 1, TOK_UNKNOWN = 2, TOK_CONTROL = 3, TOK_USER_DEFINED = 4, TOK_UNUSED = 5, TOK_BYTE = 6 };

enum ChatFormat { CHAT_CHATML = 0, CHAT_LLAMA2 = 1, CHAT_MISTRAL = 2, CHAT_ZEPHYR = 3, CHAT_LLAMA3 = 4 };

class Tokenizer {
   public:
    explicit Tokenizer(const GgufFile &f);

    // add_special: prepend BOS when the vocabulary asks for it (tokenizer.ggml.add_bos_token, default on);
    // parse_special: control / user-defined token strings in the text become their single ids
    std::vector<int32_t> encode(const std::string &text, bool add_special, bool parse_special) const;
    // render_special: emit the strings of control tokens too; strip_space: drop the one leading space the
    // SentencePiece prefix adds to the start of a text
    std::string decode(const int32_t *ids, int32_t n, bool render_special, bool strip_space) const;
    std::string piece_text(int32_t id, bool render_special) const;

    // the {system,user} message list of inference.rs:363-376 rendered for generation: the file's own template
    // (jinja_lite.h) when it can be evaluated, else the format recognised by marker strings
    std::string chat_prompt(const std::string &system, const std::string &user) const;
    // messages through tokenizer.chat_template only; false + *err when there is none or it cannot be evaluated
    bool apply_template(const std::vector<std::pair<std::string, std::string>> &messages, bool add_generation_prompt,
                        std::string *out, std::string *err) const;
    const std::string &chat_template() const { return template_; }
    ChatFormat chat_format() const { return format_; }
    // end-of-generation tokens: eos plus the turn terminators present in the vocabulary
    bool is_eog(int32_t id) const;

    int32_t n_tokens() const { return (int32_t)pieces_.size(); }
    int32_t bos() const { return bos_; }
    int32_t eos() const { return eos_; }
    bool add_bos() const { return add_bos_; }
    const std::string &model() const { return model_; }
    const std::string &piece(int32_t id) const { return pieces_[id]; }
    int32_t type(int32_t id) const { return (size_t)id < types_.size() ? types_[id] : TOK_NORMAL; }
    float score(int32_t id) const { return (size_t)id < scores_.size() ? scores_[id] : 0.f; }

   private:
    void encode_spm(const std::string &text, std::vector<int32_t> &out) const;
    void emit_symbol(const std::string &s, std::vector<int32_t> &out) const;
    void encode_bpe(const std::string &text, std::vector<int32_t> &out) const;
    void bpe_word(const std::string &mapped, std::vector<int32_t> &out) const;

    std::string model_;
    std::vector<std::string> pieces_;
    std::vector<int32_t> types_;
    std::vector<float> scores_;
    std::unordered_map<std::string, int32_t> index_;
    std::vector<int32_t> specials_;  // control + user-defined ids, longest piece first
    std::vector<int32_t> eog_;
    int32_t byte_ids_[256];
    int32_t bos_ = -1, eos_ = -1, unk_ = 0;
    bool add_bos_ = true, add_space_prefix_ = true;
    // byte-level BPE
    bool bpe_ = false, ignore_merges_ = false;
    int pre_ = 0;  // 0: GPT-2 pattern, 1: Llama-3 pattern
    std::unordered_map<std::string, int32_t> merge_rank_;  // "left right" -> rank
    ChatFormat format_ = CHAT_CHATML;
    std::string template_;  // tokenizer.chat_template
};

}  // namespace aios
