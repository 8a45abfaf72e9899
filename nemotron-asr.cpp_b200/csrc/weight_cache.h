// weight_cache.h — the native weight file: the device-ready form of the big operands as build_model lays
// them out (fp16 [N][K] matrices with q|k|v stacked, GLU rows interleaved, pre_encode.out columns permuted, the [Wh|Wh|Wl]
// split-precision triples, the decode kernel's hi/lo pairs), keyed by name, so that a later load is mmap + one cudaMemcpy per
// tensor instead of dequantise + repack on the host. The reference has no counterpart (it mmaps the GGUF and lets ggml
// dequantise on the fly, src/nemo-ggml.cpp:83-442); its converter scripts/convert_to_gguf.py:322-336 is where on-disk formats
// are chosen.
// File: "NB200WC2" | u64 source GGUF size | u64 source mtime (ns) | u64 n_entries | u64 index offset | blobs (256-byte aligned)
//       | index: per entry u32 key length, key bytes, u64 offset, u64 bytes. Little-endian, same-machine use (a cache, not an
//       interchange format). A cache whose source stamp or version differs is ignored and rewritten.
#pragma once
#include <cstdint>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

namespace nb {

struct WeightCacheBlob {
    const uint8_t* data = nullptr;
    uint64_t bytes = 0;
};

class WeightCacheReader {
   public:
    ~WeightCacheReader();
    // false (and closed) if the file is missing, malformed or was made from another GGUF
    bool open(const std::string& path, uint64_t src_size, uint64_t src_mtime_ns);
    const WeightCacheBlob* find(const std::string& key, uint64_t expect_bytes) const;
    size_t entries() const { return index_.size(); }

   private:
    void* map_ = nullptr;
    size_t len_ = 0;
    std::map<std::string, WeightCacheBlob> index_;
};

class WeightCacheWriter {
   public:
    ~WeightCacheWriter();
    void open(const std::string& path, uint64_t src_size, uint64_t src_mtime_ns);  // writes to path + ".tmp<pid>", throws nb::Error
    void add(const std::string& key, const void* data, uint64_t bytes);
    void commit();  // index + header, then rename over `path`
    bool active() const { return f_ != nullptr; }

   private:
    struct Entry { std::string key; uint64_t off, bytes; };
    FILE* f_ = nullptr;
    std::string path_, tmp_;
    uint64_t pos_ = 0;
    std::vector<Entry> entries_;
};

void file_stamp(const std::string& path, uint64_t& size, uint64_t& mtime_ns);  // throws nb::Error

}  // namespace nb
