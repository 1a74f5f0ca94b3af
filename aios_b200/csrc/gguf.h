// gguf.h — read-only GGUF v2/v3 container parser over an mmap (host side, C++17).
// Replaces the GGUF open/parse/mmap stage llama-server performs for the reference
// (SURVEY.md §8a D1).  Format per gguf-py gguf_reader.py:132-186, :310-368 and
// constants.py:10-12, :4157-4170.
#pragma once
#include <stdint.h>

#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace aios {

struct GgufError : std::runtime_error {
    int code;  // AIOS_E_IO or AIOS_E_FORMAT
    GgufError(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

struct GgufTensor {
    std::string name;
    int32_t type = 0;
    int32_t n_dims = 0;
    int64_t ne[4] = {1, 1, 1, 1};  // ne[0] innermost
    uint64_t offset = 0;           // relative to the data section
    uint64_t nbytes = 0;
    const uint8_t *data = nullptr;  // into the mmap
};

struct GgufValue {
    enum Kind { NONE, UINT, INT, FLOAT, BOOL, STRING, ARRAY } kind = NONE;
    uint64_t u = 0;
    int64_t i = 0;
    double f = 0;
    std::string s;
    // arrays: element type + raw location (strings are materialised lazily by the tokenizer)
    uint32_t arr_type = 0;
    uint64_t arr_n = 0;
    const uint8_t *arr_data = nullptr;
};

class GgufFile {
   public:
    explicit GgufFile(const std::string &path);
    ~GgufFile();
    GgufFile(const GgufFile &) = delete;
    GgufFile &operator=(const GgufFile &) = delete;

    uint32_t version() const { return version_; }
    uint64_t file_bytes() const { return size_; }
    uint32_t alignment() const { return alignment_; }
    uint64_t data_offset() const { return data_off_; }
    const std::vector<GgufTensor> &tensors() const { return tensors_; }
    const GgufTensor *find(const std::string &name) const;
    const GgufTensor &require(const std::string &name) const;

    bool has(const std::string &key) const { return kv_.count(key) != 0; }
    const GgufValue *get(const std::string &key) const;
    uint64_t get_u(const std::string &key, uint64_t dflt) const;
    double get_f(const std::string &key, double dflt) const;
    std::string get_s(const std::string &key, const std::string &dflt) const;
    // string array element access (tokenizer.ggml.tokens)
    std::vector<std::string> get_str_array(const std::string &key) const;
    std::vector<int32_t> get_i32_array(const std::string &key) const;
    std::vector<float> get_f32_array(const std::string &key) const;

   private:
    void parse();
    int fd_ = -1;
    const uint8_t *base_ = nullptr;
    uint64_t size_ = 0;
    uint32_t version_ = 0, alignment_ = 32;
    uint64_t data_off_ = 0;
    std::vector<GgufTensor> tensors_;
    std::map<std::string, size_t> index_;
    std::map<std::string, GgufValue> kv_;
};

}  // namespace aios
