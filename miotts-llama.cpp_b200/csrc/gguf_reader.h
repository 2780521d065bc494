// Minimal GGUF v2/v3 reader (file format: third_party/llama.cpp/ggml/include/gguf.h:1-31 of the
// reference).  Replaces the reference's use of gguf_init_from_file(no_alloc) + the per-tensor reads of
// preload_all_tensor_data (src/miocodec.cpp:92-135).  Host-only, no CUDA, no ggml.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace mio {

enum gguf_dtype : uint32_t { GT_F32 = 0, GT_F16 = 1, GT_I32 = 26, GT_BF16 = 30 };

struct gguf_tensor {
    std::string name;
    uint32_t    n_dims = 0;
    int64_t     ne[4]  = {1, 1, 1, 1};   // ggml order: ne[0] is the contiguous axis
    uint32_t    dtype  = 0;
    uint64_t    offset = 0;              // relative to data section
    int64_t nelements() const { return ne[0] * ne[1] * ne[2] * ne[3]; }
    size_t  nbytes() const;              // 0 for unsupported dtypes
};

struct gguf_file {
    std::string path;
    uint32_t version = 0;
    uint64_t data_offset = 0;
    uint32_t alignment = 32;
    std::map<std::string, uint32_t>    kv_u32;   // u8/u16/u32/i8/i16/i32/bool/u64/i64 (narrowed)
    std::map<std::string, float>       kv_f32;   // f32/f64
    std::map<std::string, std::string> kv_str;
    std::vector<gguf_tensor> tensors;            // file order (load_voice_embedding takes tensors[0])
    std::map<std::string, size_t> index;

    // returns empty string on success, otherwise the error text
    std::string open(const std::string & p);
    const gguf_tensor * find(const std::string & name) const;
    int   u32_or(const char * key, int def) const;       // gguf_get_val_u32_or, src/miocodec.cpp:149-152
    float f32_or(const char * key, float def) const;     // gguf_get_val_f32_or, src/miocodec.cpp:154-157
    // read raw bytes of one tensor; false on short read
    bool read_raw(const gguf_tensor & t, void * dst, size_t nbytes) const;
    // read a tensor as float32 whatever its storage type (F32/F16/BF16); false if unsupported
    bool read_f32(const gguf_tensor & t, std::vector<float> & out) const;
};

float fp16_bits_to_float(uint16_t h);

}  // namespace mio
