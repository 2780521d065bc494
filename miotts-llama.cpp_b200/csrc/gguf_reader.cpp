#include "gguf_reader.h"

#include <algorithm>
#include <cstdio>
#include <cstring>

namespace mio {

namespace {
enum kv_type : uint32_t {
    KV_U8 = 0, KV_I8 = 1, KV_U16 = 2, KV_I16 = 3, KV_U32 = 4, KV_I32 = 5, KV_F32 = 6, KV_BOOL = 7,
    KV_STR = 8, KV_ARR = 9, KV_U64 = 10, KV_I64 = 11, KV_F64 = 12
};

struct reader {
    FILE * f = nullptr;
    bool ok = true;
    ~reader() { if (f) fclose(f); }
    template <class T> T get() {
        T v{};
        if (ok && fread(&v, sizeof(T), 1, f) != 1) ok = false;
        return v;
    }
    std::string str() {
        uint64_t n = get<uint64_t>();
        if (!ok || n > (1u << 24)) { ok = false; return {}; }
        std::string s(n, '\0');
        if (n && fread(&s[0], 1, n, f) != n) ok = false;
        return s;
    }
    void skip(uint64_t n) { if (ok && fseek(f, (long) n, SEEK_CUR) != 0) ok = false; }
};

size_t scalar_size(uint32_t t) {
    switch (t) {
        case KV_U8: case KV_I8: case KV_BOOL: return 1;
        case KV_U16: case KV_I16: return 2;
        case KV_U32: case KV_I32: case KV_F32: return 4;
        case KV_U64: case KV_I64: case KV_F64: return 8;
        default: return 0;
    }
}
}  // namespace

size_t gguf_tensor::nbytes() const {
    size_t es = 0;
    switch (dtype) {
        case GT_F32: case GT_I32: es = 4; break;
        case GT_F16: case GT_BF16: es = 2; break;
        default: return 0;
    }
    return (size_t) nelements() * es;
}

float fp16_bits_to_float(uint16_t h) {
    uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1f;
    uint32_t man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else {  // subnormal: normalise
            int e = -1;
            do { man <<= 1; e++; } while (!(man & 0x400u));
            man &= 0x3ffu;
            bits = sign | ((uint32_t) (127 - 15 - e) << 23) | (man << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | (man << 13);
    } else {
        bits = sign | ((exp + 112) << 23) | (man << 13);
    }
    float out;
    memcpy(&out, &bits, 4);
    return out;
}

std::string gguf_file::open(const std::string & p) {
    path = p;
    reader r;
    r.f = fopen(p.c_str(), "rb");
    if (!r.f) return "cannot open " + p;
    char magic[4] = {0, 0, 0, 0};
    if (fread(magic, 1, 4, r.f) != 4 || memcmp(magic, "GGUF", 4) != 0) return "not a GGUF file: " + p;
    version = r.get<uint32_t>();
    if (version < 2 || version > 3) return "unsupported GGUF version in " + p;
    uint64_t n_tensors = r.get<uint64_t>();
    uint64_t n_kv = r.get<uint64_t>();
    if (!r.ok || n_tensors > (1u << 20) || n_kv > (1u << 20)) return "corrupt GGUF header: " + p;

    for (uint64_t i = 0; i < n_kv && r.ok; i++) {
        std::string key = r.str();
        uint32_t t = r.get<uint32_t>();
        switch (t) {
            case KV_U8:   kv_u32[key] = r.get<uint8_t>(); break;
            case KV_I8:   kv_u32[key] = (uint32_t) (int32_t) r.get<int8_t>(); break;
            case KV_U16:  kv_u32[key] = r.get<uint16_t>(); break;
            case KV_I16:  kv_u32[key] = (uint32_t) (int32_t) r.get<int16_t>(); break;
            case KV_U32:  kv_u32[key] = r.get<uint32_t>(); break;
            case KV_I32:  kv_u32[key] = (uint32_t) r.get<int32_t>(); break;
            case KV_BOOL: kv_u32[key] = r.get<uint8_t>(); break;
            case KV_U64:  kv_u32[key] = (uint32_t) r.get<uint64_t>(); break;
            case KV_I64:  kv_u32[key] = (uint32_t) r.get<int64_t>(); break;
            case KV_F32:  kv_f32[key] = r.get<float>(); break;
            case KV_F64:  kv_f32[key] = (float) r.get<double>(); break;
            case KV_STR:  kv_str[key] = r.str(); break;
            case KV_ARR: {
                uint32_t et = r.get<uint32_t>();
                uint64_t n = r.get<uint64_t>();
                if (et == KV_STR) {
                    for (uint64_t j = 0; j < n && r.ok; j++) (void) r.str();
                } else {
                    size_t es = scalar_size(et);
                    if (!es) r.ok = false;
                    r.skip(n * es);
                }
                break;
            }
            default: r.ok = false;
        }
    }
    if (!r.ok) return "corrupt GGUF key/value section: " + p;
    auto al = kv_u32.find("general.alignment");
    if (al != kv_u32.end() && al->second) alignment = al->second;

    tensors.resize(n_tensors);
    for (uint64_t i = 0; i < n_tensors && r.ok; i++) {
        gguf_tensor & t = tensors[i];
        t.name = r.str();
        t.n_dims = r.get<uint32_t>();
        if (t.n_dims > 4) { r.ok = false; break; }
        for (uint32_t d = 0; d < t.n_dims; d++) t.ne[d] = (int64_t) r.get<uint64_t>();
        t.dtype = r.get<uint32_t>();
        t.offset = r.get<uint64_t>();
        index[t.name] = i;
    }
    if (!r.ok) return "corrupt GGUF tensor index: " + p;
    long pos = ftell(r.f);
    data_offset = ((uint64_t) pos + alignment - 1) / alignment * alignment;
    // every tensor must lie inside the file: sizes and offsets come from the file and are used for allocations and
    // seeks later (a negative / huge extent must fail here with an error code, not as bad_alloc across the C ABI)
    if (fseek(r.f, 0, SEEK_END) != 0) return "corrupt GGUF file: " + p;
    const uint64_t file_size = (uint64_t) ftell(r.f);
    for (const gguf_tensor & t : tensors) {
        uint64_t n = 1;
        for (uint32_t d = 0; d < 4; d++) {
            if (t.ne[d] < 0 || (t.ne[d] > 0 && n > (1ull << 40) / (uint64_t) t.ne[d])) return "corrupt GGUF tensor extent (" + t.name + "): " + p;
            n *= (uint64_t) t.ne[d];
        }
        const uint64_t bytes = (uint64_t) t.nbytes();     // 0 for dtypes this reader does not handle (rejected when read)
        if (t.offset > file_size || data_offset > file_size || bytes > file_size - std::min(file_size, data_offset + t.offset))
            return "corrupt GGUF tensor data range (" + t.name + "): " + p;
    }
    return {};
}

const gguf_tensor * gguf_file::find(const std::string & name) const {
    auto it = index.find(name);
    return it == index.end() ? nullptr : &tensors[it->second];
}

int gguf_file::u32_or(const char * key, int def) const {
    auto it = kv_u32.find(key);
    return it == kv_u32.end() ? def : (int) it->second;
}

float gguf_file::f32_or(const char * key, float def) const {
    auto it = kv_f32.find(key);
    return it == kv_f32.end() ? def : it->second;
}

bool gguf_file::read_raw(const gguf_tensor & t, void * dst, size_t nbytes) const {
    FILE * f = fopen(path.c_str(), "rb");
    if (!f) return false;
    bool ok = fseek(f, (long) (data_offset + t.offset), SEEK_SET) == 0 && fread(dst, 1, nbytes, f) == nbytes;
    fclose(f);
    return ok;
}

bool gguf_file::read_f32(const gguf_tensor & t, std::vector<float> & out) const {
    const size_t n = (size_t) t.nelements();
    out.resize(n);
    if (t.dtype == GT_F32) return read_raw(t, out.data(), n * 4);
    if (t.dtype == GT_F16 || t.dtype == GT_BF16) {
        std::vector<uint16_t> tmp(n);
        if (!read_raw(t, tmp.data(), n * 2)) return false;
        if (t.dtype == GT_F16) {
            for (size_t i = 0; i < n; i++) out[i] = fp16_bits_to_float(tmp[i]);
        } else {
            for (size_t i = 0; i < n; i++) {
                uint32_t b = (uint32_t) tmp[i] << 16;
                memcpy(&out[i], &b, 4);
            }
        }
        return true;
    }
    return false;
}

}  // namespace mio
