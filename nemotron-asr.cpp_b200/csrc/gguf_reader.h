// gguf_reader.h — minimal GGUF v3 reader (host). Format: the reference's writer scripts/convert_to_gguf.py:491-536;
// keys and tensor names consumed: src/nemo-ggml.cpp:108-182, 296-398.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace nb {

enum GgmlType : int32_t { GGML_F32 = 0, GGML_F16 = 1, GGML_Q4_0 = 2, GGML_Q8_0 = 8 };

struct GgufTensor {
    std::string name;
    std::vector<int64_t> dims;  // fastest first (ggml order); PyTorch shape is the reverse
    int32_t type = 0;
    const uint8_t* data = nullptr;
    size_t nbytes = 0;
    int64_t nelements() const {
        int64_t n = 1;
        for (auto d : dims) n *= d;
        return n;
    }
};

struct GgufFile {
    std::map<std::string, uint32_t> u32;                      // scalar integer KVs
    std::map<std::string, std::string> str;                   // string KVs
    std::map<std::string, std::vector<std::string>> str_arr;  // string-array KVs
    std::map<std::string, std::vector<int32_t>> i32_arr;
    std::map<std::string, GgufTensor> tensors;

    ~GgufFile();
    void open(const std::string& path);  // throws nb::Error
    const GgufTensor& tensor(const std::string& name) const;
    bool has(const std::string& name) const { return tensors.count(name) != 0; }

   private:
    void* map_ = nullptr;
    size_t map_len_ = 0;
};

// Dequantise any supported tensor to fp32 (Q8_0: ggml/src/ggml-quants.c:553-567; Q4_0: :455-476; F16).
void dequantize_to_f32(const GgufTensor& t, float* out);

}  // namespace nb
