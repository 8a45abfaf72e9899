#include "gguf_reader.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstring>
#include <stdexcept>

#include <cuda_fp16.h>

namespace nb {

namespace {
struct Cursor {
    const uint8_t* p;
    const uint8_t* end;
    // every length read from the file is compared with the bytes that remain (never added to a pointer first: a crafted
    // 64-bit length would wrap)
    size_t left() const { return (size_t)(end - p); }
    template <typename T>
    T get() {
        if (sizeof(T) > left()) throw std::runtime_error("GGUF: truncated file");
        T v;
        memcpy(&v, p, sizeof(T));
        p += sizeof(T);
        return v;
    }
    std::string str() {
        uint64_t n = get<uint64_t>();
        if (n > left()) throw std::runtime_error("GGUF: truncated string");
        std::string s((const char*)p, (size_t)n);
        p += n;
        return s;
    }
    void skip(uint64_t elem, uint64_t cnt) {
        if (elem != 0 && cnt > left() / elem) throw std::runtime_error("GGUF: truncated file");
        p += elem * cnt;
    }
};

size_t scalar_size(uint32_t t) {
    switch (t) {
        case 0: case 1: case 7: return 1;
        case 2: case 3: return 2;
        case 4: case 5: case 6: return 4;
        case 10: case 11: case 12: return 8;
        default: throw std::runtime_error("GGUF: unknown value type");
    }
}

constexpr int64_t MAX_ELEMENTS = (int64_t)1 << 40;  // far above any tensor of this model family; keeps n * bytes inside size_t

size_t type_bytes(int32_t type, int64_t n, const std::string& name) {
    if (n < 0 || n > MAX_ELEMENTS) throw std::runtime_error("GGUF: element count out of range: " + name);
    switch (type) {
        case GGML_F32: return (size_t)n * 4;
        case GGML_F16: return (size_t)n * 2;
        case GGML_Q8_0:
        case GGML_Q4_0:
            if (n % 32 != 0) throw std::runtime_error("GGUF: quantised tensor whose element count is not a multiple of 32: " + name);
            return (size_t)(n / 32) * (type == GGML_Q8_0 ? 34 : 18);
        default: throw std::runtime_error("GGUF: unsupported tensor type " + std::to_string(type) + " (" + name + ")");
    }
}

// product of the dimensions with overflow / sign checks (GgufTensor::nelements() trusts what was validated here)
int64_t checked_elements(const std::vector<int64_t>& dims, const std::string& name) {
    int64_t n = 1;
    for (int64_t d : dims) {
        if (d < 0) throw std::runtime_error("GGUF: negative dimension in " + name);
        if (d != 0 && n > MAX_ELEMENTS / d) throw std::runtime_error("GGUF: element count out of range: " + name);
        n *= d;
    }
    return n;
}
}  // namespace

GgufFile::~GgufFile() {
    if (map_) munmap(map_, map_len_);
}

void GgufFile::open(const std::string& path) {
    int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("GGUF: cannot open " + path);
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size <= 0) {
        ::close(fd);
        throw std::runtime_error("GGUF: cannot stat " + path);
    }
    map_len_ = (size_t)st.st_size;
    map_ = mmap(nullptr, map_len_, PROT_READ, MAP_PRIVATE, fd, 0);
    ::close(fd);
    if (map_ == MAP_FAILED) {
        map_ = nullptr;
        throw std::runtime_error("GGUF: mmap failed for " + path);
    }
    Cursor c{(const uint8_t*)map_, (const uint8_t*)map_ + map_len_};
    if (map_len_ < 24 || memcmp(c.p, "GGUF", 4) != 0) throw std::runtime_error("GGUF: bad magic in " + path);
    c.p += 4;
    uint32_t version = c.get<uint32_t>();
    if (version != 3 && version != 2) throw std::runtime_error("GGUF: unsupported version");
    int64_t n_tensors = c.get<int64_t>();
    int64_t n_kv = c.get<int64_t>();
    // every KV / tensor record takes at least 12 bytes of the file: counts beyond that are garbage
    if (n_tensors < 0 || n_kv < 0 || (uint64_t)n_tensors > map_len_ / 12 || (uint64_t)n_kv > map_len_ / 12)
        throw std::runtime_error("GGUF: implausible tensor / KV count");
    uint32_t alignment = 32;
    for (int64_t i = 0; i < n_kv; i++) {
        std::string key = c.str();
        uint32_t t = c.get<uint32_t>();
        if (t == 8) {
            str[key] = c.str();
        } else if (t == 9) {
            uint32_t et = c.get<uint32_t>();
            uint64_t cnt = c.get<uint64_t>();
            if (et == 8) {
                if (cnt > c.left() / 8) throw std::runtime_error("GGUF: truncated string array");  // 8 bytes of length per entry at least
                auto& v = str_arr[key];
                v.reserve((size_t)cnt);
                for (uint64_t j = 0; j < cnt; j++) v.push_back(c.str());
            } else if (et == 5) {
                if (cnt > c.left() / 4) throw std::runtime_error("GGUF: truncated i32 array");
                auto& v = i32_arr[key];
                v.reserve((size_t)cnt);
                for (uint64_t j = 0; j < cnt; j++) v.push_back(c.get<int32_t>());
            } else {
                c.skip(scalar_size(et), cnt);
            }
        } else if (t == 4 || t == 5) {
            u32[key] = c.get<uint32_t>();
        } else {
            c.skip(scalar_size(t), 1);
        }
    }
    if (u32.count("general.alignment")) alignment = u32["general.alignment"];
    if (alignment == 0 || alignment > (1u << 20) || (alignment & (alignment - 1)) != 0) throw std::runtime_error("GGUF: bad general.alignment");
    struct Info {
        std::string name;
        std::vector<int64_t> dims;
        int32_t type;
        uint64_t off;
    };
    std::vector<Info> infos;
    for (int64_t i = 0; i < n_tensors; i++) {
        Info in;
        in.name = c.str();
        uint32_t nd = c.get<uint32_t>();
        if (nd > 4) throw std::runtime_error("GGUF: tensor with more than 4 dimensions: " + in.name);  // GGML_MAX_DIMS
        for (uint32_t d = 0; d < nd; d++) in.dims.push_back(c.get<int64_t>());
        in.type = c.get<int32_t>();
        in.off = c.get<uint64_t>();
        infos.push_back(std::move(in));
    }
    size_t pos = (size_t)(c.p - (const uint8_t*)map_);
    size_t start = (pos + alignment - 1) / alignment * alignment;
    for (auto& in : infos) {
        GgufTensor t;
        t.name = in.name;
        t.dims = in.dims;
        t.type = in.type;
        t.nbytes = type_bytes(in.type, checked_elements(in.dims, in.name), in.name);
        if (start > map_len_ || in.off > map_len_ - start || t.nbytes > map_len_ - start - in.off)
            throw std::runtime_error("GGUF: tensor data out of range: " + in.name);
        t.data = (const uint8_t*)map_ + start + in.off;
        tensors[in.name] = std::move(t);
    }
}

const GgufTensor& GgufFile::tensor(const std::string& name) const {
    auto it = tensors.find(name);
    if (it == tensors.end()) throw std::runtime_error("GGUF: missing tensor: " + name);
    return it->second;
}

void dequantize_to_f32(const GgufTensor& t, float* out) {
    int64_t n = t.nelements();
    if (t.type == GGML_F32) {
        memcpy(out, t.data, (size_t)n * 4);
    } else if (t.type == GGML_F16) {
        const __half* h = (const __half*)t.data;
        for (int64_t i = 0; i < n; i++) out[i] = __half2float(h[i]);
    } else if (t.type == GGML_Q8_0) {
        const uint8_t* p = t.data;
        for (int64_t b = 0; b < n / 32; b++, p += 34) {
            __half dh;
            memcpy(&dh, p, 2);
            float d = __half2float(dh);
            const int8_t* q = (const int8_t*)(p + 2);
            for (int j = 0; j < 32; j++) out[b * 32 + j] = (float)q[j] * d;
        }
    } else if (t.type == GGML_Q4_0) {
        const uint8_t* p = t.data;
        for (int64_t b = 0; b < n / 32; b++, p += 18) {
            __half dh;
            memcpy(&dh, p, 2);
            float d = __half2float(dh);
            for (int j = 0; j < 16; j++) {
                out[b * 32 + j] = (float)((int)(p[2 + j] & 0x0F) - 8) * d;
                out[b * 32 + j + 16] = (float)((int)(p[2 + j] >> 4) - 8) * d;
            }
        }
    } else {
        throw std::runtime_error("dequantize: unsupported type");
    }
}

}  // namespace nb
