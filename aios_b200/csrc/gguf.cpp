// gguf.cpp — see gguf.h.
#include "gguf.h"

#include <errno.h>
#include <fcntl.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/aios_b200.h"
#include "layout.cuh"

namespace aios {

namespace {
enum : uint32_t { GT_U8 = 0, GT_I8, GT_U16, GT_I16, GT_U32, GT_I32, GT_F32, GT_BOOL, GT_STR, GT_ARR, GT_U64, GT_I64, GT_F64 };

size_t scalar_size(uint32_t t) {
    switch (t) {
        case GT_U8: case GT_I8: case GT_BOOL: return 1;
        case GT_U16: case GT_I16: return 2;
        case GT_U32: case GT_I32: case GT_F32: return 4;
        case GT_U64: case GT_I64: case GT_F64: return 8;
    }
    return 0;
}

struct Cursor {
    const uint8_t *p, *end;
    template <typename T>
    T rd() {
        if ((size_t)(end - p) < sizeof(T)) throw GgufError(AIOS_E_FORMAT, "GGUF: truncated file");
        T v;
        memcpy(&v, p, sizeof(T));
        p += sizeof(T);
        return v;
    }
    std::string str() {
        const uint64_t n = rd<uint64_t>();
        if ((uint64_t)(end - p) < n) throw GgufError(AIOS_E_FORMAT, "GGUF: truncated string");
        std::string s(reinterpret_cast<const char *>(p), (size_t)n);
        p += n;
        return s;
    }
    void skip(uint64_t n) {
        if ((uint64_t)(end - p) < n) throw GgufError(AIOS_E_FORMAT, "GGUF: truncated file");
        p += n;
    }
};
}  // namespace

GgufFile::GgufFile(const std::string &path) {
    fd_ = ::open(path.c_str(), O_RDONLY);
    if (fd_ < 0) throw GgufError(AIOS_E_IO, "cannot open '" + path + "': " + strerror(errno));
    struct stat st;
    if (fstat(fd_, &st) != 0 || st.st_size < 24) {
        ::close(fd_);
        fd_ = -1;
        throw GgufError(AIOS_E_IO, "cannot stat '" + path + "' or file too small to be GGUF");
    }
    size_ = (uint64_t)st.st_size;
    void *m = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
    if (m == MAP_FAILED) {
        ::close(fd_);
        fd_ = -1;
        throw GgufError(AIOS_E_IO, "mmap failed for '" + path + "'");
    }
    base_ = static_cast<const uint8_t *>(m);
    try {
        parse();
    } catch (...) {
        munmap(const_cast<uint8_t *>(base_), size_);
        ::close(fd_);
        base_ = nullptr;
        fd_ = -1;
        throw;
    }
}

GgufFile::~GgufFile() {
    if (base_) munmap(const_cast<uint8_t *>(base_), size_);
    if (fd_ >= 0) ::close(fd_);
}

void GgufFile::parse() {
    Cursor c{base_, base_ + size_};
    if (c.rd<uint32_t>() != 0x46554747u) throw GgufError(AIOS_E_FORMAT, "not a GGUF file (bad magic)");
    version_ = c.rd<uint32_t>();
    if (version_ < 2 || version_ > 3) throw GgufError(AIOS_E_FORMAT, "unsupported GGUF version " + std::to_string(version_));
    const uint64_t n_tensors = c.rd<uint64_t>();
    const uint64_t n_kv = c.rd<uint64_t>();
    if (n_tensors > (1u << 20) || n_kv > (1u << 20)) throw GgufError(AIOS_E_FORMAT, "GGUF: implausible counts");
    for (uint64_t i = 0; i < n_kv; i++) {
        std::string key = c.str();
        const uint32_t t = c.rd<uint32_t>();
        GgufValue v;
        switch (t) {
            case GT_U8: v.kind = GgufValue::UINT; v.u = c.rd<uint8_t>(); break;
            case GT_I8: v.kind = GgufValue::INT; v.i = c.rd<int8_t>(); break;
            case GT_U16: v.kind = GgufValue::UINT; v.u = c.rd<uint16_t>(); break;
            case GT_I16: v.kind = GgufValue::INT; v.i = c.rd<int16_t>(); break;
            case GT_U32: v.kind = GgufValue::UINT; v.u = c.rd<uint32_t>(); break;
            case GT_I32: v.kind = GgufValue::INT; v.i = c.rd<int32_t>(); break;
            case GT_F32: v.kind = GgufValue::FLOAT; v.f = c.rd<float>(); break;
            case GT_BOOL: v.kind = GgufValue::BOOL; v.u = c.rd<uint8_t>(); break;
            case GT_STR: v.kind = GgufValue::STRING; v.s = c.str(); break;
            case GT_U64: v.kind = GgufValue::UINT; v.u = c.rd<uint64_t>(); break;
            case GT_I64: v.kind = GgufValue::INT; v.i = c.rd<int64_t>(); break;
            case GT_F64: v.kind = GgufValue::FLOAT; v.f = c.rd<double>(); break;
            case GT_ARR: {
                v.kind = GgufValue::ARRAY;
                v.arr_type = c.rd<uint32_t>();
                v.arr_n = c.rd<uint64_t>();
                v.arr_data = c.p;
                if (v.arr_type == GT_STR) {
                    for (uint64_t j = 0; j < v.arr_n; j++) c.skip(c.rd<uint64_t>());
                } else if (v.arr_type == GT_ARR) {
                    throw GgufError(AIOS_E_FORMAT, "GGUF: nested arrays are not supported (key " + key + ")");
                } else {
                    const size_t es = scalar_size(v.arr_type);
                    if (!es) throw GgufError(AIOS_E_FORMAT, "GGUF: bad array element type");
                    c.skip(v.arr_n * es);
                }
                break;
            }
            default: throw GgufError(AIOS_E_FORMAT, "GGUF: unknown value type " + std::to_string(t) + " for key " + key);
        }
        if (v.kind == GgufValue::INT) v.u = (uint64_t)v.i;
        if (v.kind == GgufValue::UINT || v.kind == GgufValue::BOOL) v.i = (int64_t)v.u;
        kv_[key] = std::move(v);
    }
    alignment_ = (uint32_t)get_u("general.alignment", 32);
    if (alignment_ == 0 || (alignment_ & (alignment_ - 1))) throw GgufError(AIOS_E_FORMAT, "GGUF: alignment is not a power of two");
    tensors_.reserve((size_t)n_tensors);
    for (uint64_t i = 0; i < n_tensors; i++) {
        GgufTensor t;
        t.name = c.str();
        t.n_dims = (int32_t)c.rd<uint32_t>();
        if (t.n_dims < 1 || t.n_dims > 4) throw GgufError(AIOS_E_FORMAT, "GGUF: tensor '" + t.name + "' has bad rank");
        for (int d = 0; d < t.n_dims; d++) t.ne[d] = (int64_t)c.rd<uint64_t>();
        t.type = (int32_t)c.rd<uint32_t>();
        t.offset = c.rd<uint64_t>();
        const int be = type_block_elems(t.type), bb = type_block_bytes(t.type);
        if (bb == 0) throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "' has unsupported ggml type " + std::to_string(t.type));
        if (t.ne[0] % be) throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "': row length not a multiple of the block size");
        uint64_t n = 1;
        for (int d = 0; d < t.n_dims; d++) n *= (uint64_t)t.ne[d];
        t.nbytes = n / (uint64_t)be * (uint64_t)bb;
        index_[t.name] = tensors_.size();
        tensors_.push_back(std::move(t));
    }
    const uint64_t hdr = (uint64_t)(c.p - base_);
    data_off_ = (hdr + alignment_ - 1) / alignment_ * alignment_;
    for (auto &t : tensors_) {
        if (t.offset % alignment_) throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "' is misaligned");
        if (data_off_ + t.offset + t.nbytes > size_) throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "' runs past the end of the file");
        t.data = base_ + data_off_ + t.offset;
    }
}

const GgufTensor *GgufFile::find(const std::string &name) const {
    auto it = index_.find(name);
    return it == index_.end() ? nullptr : &tensors_[it->second];
}
const GgufTensor &GgufFile::require(const std::string &name) const {
    const GgufTensor *t = find(name);
    if (!t) throw GgufError(AIOS_E_FORMAT, "GGUF: missing tensor '" + name + "'");
    return *t;
}
const GgufValue *GgufFile::get(const std::string &key) const {
    auto it = kv_.find(key);
    return it == kv_.end() ? nullptr : &it->second;
}
uint64_t GgufFile::get_u(const std::string &key, uint64_t dflt) const {
    const GgufValue *v = get(key);
    if (!v) return dflt;
    if (v->kind == GgufValue::UINT || v->kind == GgufValue::INT || v->kind == GgufValue::BOOL) return v->u;
    if (v->kind == GgufValue::FLOAT) return (uint64_t)v->f;
    return dflt;
}
double GgufFile::get_f(const std::string &key, double dflt) const {
    const GgufValue *v = get(key);
    if (!v) return dflt;
    if (v->kind == GgufValue::FLOAT) return v->f;
    if (v->kind == GgufValue::UINT) return (double)v->u;
    if (v->kind == GgufValue::INT) return (double)v->i;
    return dflt;
}
std::string GgufFile::get_s(const std::string &key, const std::string &dflt) const {
    const GgufValue *v = get(key);
    return (v && v->kind == GgufValue::STRING) ? v->s : dflt;
}
std::vector<std::string> GgufFile::get_str_array(const std::string &key) const {
    std::vector<std::string> out;
    const GgufValue *v = get(key);
    if (!v || v->kind != GgufValue::ARRAY || v->arr_type != GT_STR) return out;
    Cursor c{v->arr_data, base_ + size_};
    out.reserve((size_t)v->arr_n);
    for (uint64_t i = 0; i < v->arr_n; i++) out.push_back(c.str());
    return out;
}
std::vector<int32_t> GgufFile::get_i32_array(const std::string &key) const {
    std::vector<int32_t> out;
    const GgufValue *v = get(key);
    if (!v || v->kind != GgufValue::ARRAY || (v->arr_type != GT_I32 && v->arr_type != GT_U32)) return out;
    out.resize((size_t)v->arr_n);
    memcpy(out.data(), v->arr_data, out.size() * 4);
    return out;
}
std::vector<float> GgufFile::get_f32_array(const std::string &key) const {
    std::vector<float> out;
    const GgufValue *v = get(key);
    if (!v || v->kind != GgufValue::ARRAY || v->arr_type != GT_F32) return out;
    out.resize((size_t)v->arr_n);
    memcpy(out.data(), v->arr_data, out.size() * 4);
    return out;
}

}  // namespace aios
