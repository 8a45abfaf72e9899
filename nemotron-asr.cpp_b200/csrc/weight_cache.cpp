// weight_cache.cpp — see weight_cache.h
#include "weight_cache.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstring>

#include "common.cuh"

namespace nb {

namespace {
constexpr char MAGIC[8] = {'N', 'B', '2', '0', '0', 'W', 'C', '2'};
constexpr uint64_t HEADER = 8 + 4 * 8, ALIGN = 256, MAX_ENTRIES = 1 << 20, MAX_KEY = 4096;
}  // namespace

void file_stamp(const std::string& path, uint64_t& size, uint64_t& mtime_ns) {
    struct stat st;
    if (stat(path.c_str(), &st) != 0) throw Error("cannot stat " + path);
    size = (uint64_t)st.st_size;
    mtime_ns = (uint64_t)st.st_mtim.tv_sec * 1000000000ull + (uint64_t)st.st_mtim.tv_nsec;
}

WeightCacheReader::~WeightCacheReader() {
    if (map_) munmap(map_, len_);
}

bool WeightCacheReader::open(const std::string& path, uint64_t src_size, uint64_t src_mtime_ns) {
    const int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) != 0 || (uint64_t)st.st_size < HEADER) {
        close(fd);
        return false;
    }
    len_ = (size_t)st.st_size;
    map_ = mmap(nullptr, len_, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (map_ == MAP_FAILED) {
        map_ = nullptr;
        return false;
    }
    const uint8_t* p = static_cast<const uint8_t*>(map_);
    auto fail = [&] {
        munmap(map_, len_);
        map_ = nullptr;
        index_.clear();
        return false;
    };
    uint64_t h[4];
    memcpy(h, p + 8, sizeof h);
    if (memcmp(p, MAGIC, 8) != 0 || h[0] != src_size || h[1] != src_mtime_ns) return fail();
    const uint64_t n = h[2];
    uint64_t off = h[3];
    if (n > MAX_ENTRIES || off < HEADER || off > len_) return fail();
    for (uint64_t i = 0; i < n; i++) {
        uint32_t kl;
        if (len_ - off < 4) return fail();
        memcpy(&kl, p + off, 4);
        off += 4;
        if (kl > MAX_KEY || len_ - off < (uint64_t)kl + 16) return fail();
        std::string key(reinterpret_cast<const char*>(p + off), kl);
        off += kl;
        uint64_t ob[2];
        memcpy(ob, p + off, 16);
        off += 16;
        if (ob[0] < HEADER || ob[0] > len_ || ob[1] > len_ - ob[0]) return fail();  // blob inside the file (no wrap: compared against what is left)
        index_[key] = WeightCacheBlob{p + ob[0], ob[1]};
    }
    madvise(map_, len_, MADV_SEQUENTIAL);
    return true;
}

const WeightCacheBlob* WeightCacheReader::find(const std::string& key, uint64_t expect_bytes) const {
    auto it = index_.find(key);
    if (it == index_.end() || it->second.bytes != expect_bytes) return nullptr;
    return &it->second;
}

WeightCacheWriter::~WeightCacheWriter() {
    if (f_) {  // not committed: drop the partial file
        fclose(f_);
        unlink(tmp_.c_str());
    }
}

void WeightCacheWriter::open(const std::string& path, uint64_t src_size, uint64_t src_mtime_ns) {
    path_ = path;
    tmp_ = path + ".tmp" + std::to_string((long long)getpid());
    f_ = fopen(tmp_.c_str(), "wb");
    if (!f_) throw Error("cannot write the weight cache " + tmp_);
    uint8_t hdr[HEADER] = {0};
    memcpy(hdr, MAGIC, 8);
    const uint64_t h[4] = {src_size, src_mtime_ns, 0, 0};
    memcpy(hdr + 8, h, sizeof h);
    if (fwrite(hdr, 1, HEADER, f_) != HEADER) throw Error("weight cache: write failed");
    pos_ = HEADER;
}

void WeightCacheWriter::add(const std::string& key, const void* data, uint64_t bytes) {
    if (!f_) return;
    static const uint8_t zeros[ALIGN] = {0};
    const uint64_t pad = (ALIGN - pos_ % ALIGN) % ALIGN;
    if ((pad && fwrite(zeros, 1, pad, f_) != pad) || (bytes && fwrite(data, 1, bytes, f_) != bytes)) {
        fclose(f_);  // disk full ...: give the cache up (commit() then does nothing), the load itself goes on
        f_ = nullptr;
        unlink(tmp_.c_str());
        return;
    }
    pos_ += pad;
    entries_.push_back({key, pos_, bytes});
    pos_ += bytes;
}

void WeightCacheWriter::commit() {
    if (!f_) return;
    const uint64_t index_off = pos_;
    for (const Entry& e : entries_) {
        const uint32_t kl = (uint32_t)e.key.size();
        const uint64_t ob[2] = {e.off, e.bytes};
        if (fwrite(&kl, 4, 1, f_) != 1 || fwrite(e.key.data(), 1, kl, f_) != kl || fwrite(ob, 8, 2, f_) != 2) throw Error("weight cache: write failed");
    }
    const uint64_t h[2] = {(uint64_t)entries_.size(), index_off};
    if (fseek(f_, 8 + 16, SEEK_SET) != 0 || fwrite(h, 8, 2, f_) != 2) throw Error("weight cache: write failed");
    if (fclose(f_) != 0) {
        f_ = nullptr;
        unlink(tmp_.c_str());
        throw Error("weight cache: close failed");
    }
    f_ = nullptr;
    if (rename(tmp_.c_str(), path_.c_str()) != 0) {
        unlink(tmp_.c_str());
        throw Error("weight cache: cannot rename to " + path_);
    }
}

}  // namespace nb
