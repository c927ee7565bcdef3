// SURVEY §8(f)-4, second half: the reference's on-disk cache blobs, so that a kNN graph or a P matrix computed by this back-end
// can be dropped into the reference's stage cache and loaded by it (and the other way round).  HOST code only (no kernels).
//
// Layouts restated from sph/utils/FileIO.{hpp,cpp} (size_t = 8 bytes, int = 4 bytes, little endian):
//   writeCompressedVecToBinary<T>        FileIO.hpp:165-291   [u64 totalBytes][u64 numItems][u64 numChunks]
//       single  (totalBytes <  INT_MAX):                      [i32 compressedSize][LZ4 block]
//       batches (totalBytes >= INT_MAX): numChunks = 2 * ceil(totalBytes / INT_MAX), items split evenly (the first `remainder`
//                                        chunks one longer, divide_into_spans, FileIO.cpp:140-155), per chunk
//                                        [i32 compressedSize][i32 originalSize][LZ4 block]
//   writeCompressedGraphToBinary         FileIO.cpp:586-618   <base>Distances.cache (compressed float), <base>Indices.cache
//                                        (compressed int64), <base>NNs.cache (raw int64 per point), <base>Symmetric.cache (1 byte)
//   writeCompressedSparseMatHDIToBinary  FileIO.cpp:620-799   serial buffer [u64 rows]{[u64 rowSize]{[u32 col][f32 val]}} compressed
//                                        as above; in batches the ROWS are split and every chunk starts with its own [u64 rows],
//                                        chunk header [u32 compressedSize][u32 originalSize]
//   loaders                              FileIO.hpp:523-640, FileIO.cpp:895-937, :1274-1460
//
// The blocks are LZ4 *block format* (lz4.h LZ4_compress_default / LZ4_decompress_safe; lz4 is a vcpkg dependency of the
// reference, absent here).  The format is published (lz4_Block_format.md): sequences of [token: literal length << 4 | match
// length - 4][extra literal-length bytes][literals][u16 offset][extra match-length bytes]; the last sequence is literals only,
// the last 5 bytes of a block are literals and the last match starts at least 12 bytes before the end.  Any valid block decodes
// with LZ4_decompress_safe, so the writer below (a greedy single-probe hash matcher) need not reproduce liblz4's byte stream;
// tests/test_cache_io.py checks both directions against the real liblz4 that pyarrow bundles (codec "lz4_raw").
#include "../../include/sphb.h"
#include "common.cuh"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace sphb {
namespace {

int64_t g_chunk_limit = INT_MAX;  // tests lower it to reach the batched layouts with small data

inline uint32_t rd32(const uint8_t* p) {
    uint32_t v;
    std::memcpy(&v, p, 4);
    return v;
}
inline int64_t lz4_bound(int64_t n) { return n + n / 255 + 16; }

void put_len(std::vector<uint8_t>& out, int64_t rest) {  // the bytes that follow a saturated 4-bit length
    while (rest >= 255) {
        out.push_back(255);
        rest -= 255;
    }
    out.push_back(static_cast<uint8_t>(rest));
}

void lz4_compress(const uint8_t* src, int64_t n, std::vector<uint8_t>& out) {
    out.clear();
    out.reserve(static_cast<size_t>(lz4_bound(n)));
    constexpr int HBITS = 16;
    std::vector<int64_t> table(size_t(1) << HBITS, -1);
    int64_t anchor = 0, ip = 0;
    const int64_t mflimit = n - 12, matchlimit = n - 5;
    int64_t misses = 0;
    while (ip <= mflimit) {
        const uint32_t seq = rd32(src + ip);
        const uint32_t h = (seq * 2654435761u) >> (32 - HBITS);
        const int64_t ref = table[h];
        table[h] = ip;
        if (ref >= 0 && ip - ref <= 65535 && rd32(src + ref) == seq) {
            int64_t mlen = 4;
            while (ip + mlen < matchlimit && src[ref + mlen] == src[ip + mlen]) ++mlen;
            const int64_t lit = ip - anchor;
            out.push_back(static_cast<uint8_t>((std::min<int64_t>(lit, 15) << 4) | std::min<int64_t>(mlen - 4, 15)));
            if (lit >= 15) put_len(out, lit - 15);
            out.insert(out.end(), src + anchor, src + ip);
            const uint16_t off = static_cast<uint16_t>(ip - ref);
            out.push_back(static_cast<uint8_t>(off & 0xff));
            out.push_back(static_cast<uint8_t>(off >> 8));
            if (mlen - 4 >= 15) put_len(out, mlen - 4 - 15);
            ip += mlen;
            anchor = ip;
            misses = 0;
        } else {
            ip += 1 + (misses++ >> 6);  // incompressible stretches are crossed with growing steps
        }
    }
    const int64_t lit = n - anchor;
    out.push_back(static_cast<uint8_t>(std::min<int64_t>(lit, 15) << 4));
    if (lit >= 15) put_len(out, lit - 15);
    out.insert(out.end(), src + anchor, src + n);
}

// LZ4_decompress_safe: returns the number of bytes written, -1 on a malformed block or when `cap` would be exceeded.
int64_t lz4_decompress(const uint8_t* src, int64_t n, uint8_t* dst, int64_t cap) {
    int64_t ip = 0, op = 0;
    while (ip < n) {
        const uint8_t token = src[ip++];
        int64_t lit = token >> 4;
        if (lit == 15) {
            uint8_t b;
            do {
                if (ip >= n) return -1;
                b = src[ip++];
                lit += b;
            } while (b == 255);
        }
        if (ip + lit > n || op + lit > cap) return -1;
        std::memcpy(dst + op, src + ip, static_cast<size_t>(lit));
        ip += lit;
        op += lit;
        if (ip >= n) break;  // the last sequence has no match
        if (ip + 2 > n) return -1;
        const int64_t off = src[ip] | (src[ip + 1] << 8);
        ip += 2;
        if (off == 0 || off > op) return -1;
        int64_t mlen = (token & 15) + 4;
        if ((token & 15) == 15) {
            uint8_t b;
            do {
                if (ip >= n) return -1;
                b = src[ip++];
                mlen += b;
            } while (b == 255);
        }
        if (op + mlen > cap) return -1;
        for (int64_t i = 0; i < mlen; ++i) dst[op + i] = dst[op + i - off];  // byte-wise: overlapping copies are the format's RLE
        op += mlen;
    }
    return op;
}

struct File {
    FILE* f = nullptr;
    File(const std::string& path, const char* mode) : f(std::fopen(path.c_str(), mode)) {}
    ~File() {
        if (f) std::fclose(f);
    }
    bool put(const void* p, size_t bytes) { return bytes == 0 || std::fwrite(p, 1, bytes, f) == bytes; }
    bool get(void* p, size_t bytes) { return bytes == 0 || std::fread(p, 1, bytes, f) == bytes; }
};

int64_t num_chunks_for(uint64_t total_bytes) {  // FileIO.hpp:213-218
    const int64_t div = static_cast<int64_t>(std::ceil(static_cast<double>(total_bytes) / static_cast<double>(g_chunk_limit)));
    return std::max<int64_t>(1, div) * 2;
}

int write_vec(const std::string& path, const uint8_t* data, int64_t count, int elem) {
    File fo(path, "wb");
    if (!fo.f) return fail(SPHB_ERR_INVALID, "cache: cannot open " + path + " for writing");
    const uint64_t total = static_cast<uint64_t>(count) * elem, items = static_cast<uint64_t>(count);
    std::vector<uint8_t> buf;
    bool ok = true;
    if (total < static_cast<uint64_t>(g_chunk_limit)) {
        const uint64_t chunks = 1;
        lz4_compress(data, static_cast<int64_t>(total), buf);
        const int32_t csize = static_cast<int32_t>(buf.size());
        ok = fo.put(&total, 8) && fo.put(&items, 8) && fo.put(&chunks, 8) && fo.put(&csize, 4) && fo.put(buf.data(), buf.size());
    } else {
        const uint64_t chunks = static_cast<uint64_t>(num_chunks_for(total));
        ok = fo.put(&total, 8) && fo.put(&items, 8) && fo.put(&chunks, 8);
        const int64_t base = count / static_cast<int64_t>(chunks), rem = count % static_cast<int64_t>(chunks);
        int64_t at = 0;
        for (uint64_t c = 0; ok && c < chunks; ++c) {
            const int64_t len = base + (static_cast<int64_t>(c) < rem ? 1 : 0);
            lz4_compress(data + at * elem, len * elem, buf);
            const int32_t csize = static_cast<int32_t>(buf.size()), osize = static_cast<int32_t>(len * elem);
            ok = fo.put(&csize, 4) && fo.put(&osize, 4) && fo.put(buf.data(), buf.size());
            at += len;
        }
    }
    return ok ? SPHB_OK : fail(SPHB_ERR_INVALID, "cache: short write to " + path);
}

// out == nullptr: only *count is filled (the caller sizes its buffer and calls again)
int read_vec(const std::string& path, uint8_t* out, int64_t cap, int elem, int64_t* count) {
    File fi(path, "rb");
    if (!fi.f) return fail(SPHB_ERR_INVALID, "cache: cannot open " + path);
    uint64_t total = 0, items = 0, chunks = 0;
    if (!fi.get(&total, 8) || !fi.get(&items, 8) || !fi.get(&chunks, 8)) return fail(SPHB_ERR_INVALID, "cache: " + path + " has no header");
    if (total != items * static_cast<uint64_t>(elem)) return fail(SPHB_ERR_INVALID, "cache: " + path + " holds elements of another size");
    *count = static_cast<int64_t>(items);
    if (!out) return SPHB_OK;
    if (cap < static_cast<int64_t>(items)) return fail(SPHB_ERR_INVALID, "cache: output capacity too small for " + path);
    std::vector<uint8_t> buf;
    if (chunks == 1) {
        int32_t csize = 0;
        if (!fi.get(&csize, 4) || csize < 0) return fail(SPHB_ERR_INVALID, "cache: " + path + " is truncated");
        buf.resize(static_cast<size_t>(csize));
        if (!fi.get(buf.data(), buf.size())) return fail(SPHB_ERR_INVALID, "cache: " + path + " is truncated");
        if (lz4_decompress(buf.data(), csize, out, static_cast<int64_t>(total)) != static_cast<int64_t>(total))
            return fail(SPHB_ERR_INVALID, "cache: " + path + " does not decompress to its stated size");
        return SPHB_OK;
    }
    int64_t at = 0;
    for (uint64_t c = 0; c < chunks; ++c) {
        int32_t csize = 0, osize = 0;
        if (!fi.get(&csize, 4) || !fi.get(&osize, 4) || csize < 0 || osize < 0) return fail(SPHB_ERR_INVALID, "cache: " + path + " is truncated");
        buf.resize(static_cast<size_t>(csize));
        if (!fi.get(buf.data(), buf.size())) return fail(SPHB_ERR_INVALID, "cache: " + path + " is truncated");
        if (at + osize > static_cast<int64_t>(total) || lz4_decompress(buf.data(), csize, out + at, osize) != osize)
            return fail(SPHB_ERR_INVALID, "cache: a chunk of " + path + " does not decompress to its stated size");
        at += osize;
    }
    return at == static_cast<int64_t>(total) ? SPHB_OK : fail(SPHB_ERR_INVALID, "cache: " + path + " is incomplete");
}

int write_raw(const std::string& path, const void* data, size_t bytes) {  // writeVecToBinary, FileIO.hpp:89-99
    File fo(path, "wb");
    if (!fo.f) return fail(SPHB_ERR_INVALID, "cache: cannot open " + path + " for writing");
    return fo.put(data, bytes) ? SPHB_OK : fail(SPHB_ERR_INVALID, "cache: short write to " + path);
}
int read_raw(const std::string& path, std::vector<uint8_t>& out) {  // loadVecFromBinary, FileIO.hpp:435-455
    File fi(path, "rb");
    if (!fi.f) return fail(SPHB_ERR_INVALID, "cache: cannot open " + path);
    std::fseek(fi.f, 0, SEEK_END);
    const long sz = std::ftell(fi.f);
    std::fseek(fi.f, 0, SEEK_SET);
    out.resize(static_cast<size_t>(std::max<long>(sz, 0)));
    return fi.get(out.data(), out.size()) ? SPHB_OK : fail(SPHB_ERR_INVALID, "cache: short read from " + path);
}

// serial form of rows [r0, r1) of a P matrix: [u64 rows]{[u64 size]{[u32 col][f32 val]}}
void serialize_rows(const int64_t* rowptr, const uint32_t* col, const float* val, int64_t r0, int64_t r1, std::vector<uint8_t>& buf) {
    const uint64_t rows = static_cast<uint64_t>(r1 - r0);
    buf.resize(static_cast<size_t>(8 + rows * 8 + static_cast<uint64_t>(rowptr[r1] - rowptr[r0]) * 8));
    uint8_t* w = buf.data();
    std::memcpy(w, &rows, 8);
    w += 8;
    for (int64_t r = r0; r < r1; ++r) {
        const uint64_t sz = static_cast<uint64_t>(rowptr[r + 1] - rowptr[r]);
        std::memcpy(w, &sz, 8);
        w += 8;
        for (int64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
            std::memcpy(w, col + e, 4);
            std::memcpy(w + 4, val + e, 4);
            w += 8;
        }
    }
}

// Parses one serial buffer; appends to (rowptr, col, val) when they are given, always advances *row / *nnz.
bool parse_rows(const uint8_t* p, int64_t bytes, int64_t* row, int64_t* nnz, int64_t cap_n, int64_t cap_nnz, int64_t* rowptr, uint32_t* col, float* val) {
    if (bytes < 8) return false;
    uint64_t rows;
    std::memcpy(&rows, p, 8);
    int64_t at = 8;
    for (uint64_t r = 0; r < rows; ++r) {
        if (at + 8 > bytes) return false;
        uint64_t sz;
        std::memcpy(&sz, p + at, 8);
        at += 8;
        if (at + static_cast<int64_t>(sz) * 8 > bytes) return false;
        if (rowptr) {
            if (*row >= cap_n || *nnz + static_cast<int64_t>(sz) > cap_nnz) return false;
            for (uint64_t e = 0; e < sz; ++e) {
                std::memcpy(col + *nnz + e, p + at + e * 8, 4);
                std::memcpy(val + *nnz + e, p + at + e * 8 + 4, 4);
            }
            rowptr[*row + 1] = *nnz + static_cast<int64_t>(sz);
        }
        at += static_cast<int64_t>(sz) * 8;
        *nnz += static_cast<int64_t>(sz);
        ++*row;
    }
    return at == bytes;
}

}  // namespace
}  // namespace sphb
using namespace sphb;

extern "C" {

SPHB_API void sphb_debug_cache_chunk_limit(int64_t bytes) { g_chunk_limit = bytes > 0 ? bytes : INT_MAX; }

SPHB_API int64_t sphb_lz4_compress_bound(int64_t n) { return lz4_bound(n); }
SPHB_API int64_t sphb_lz4_compress(const void* src, int64_t n, void* dst, int64_t cap) {
    if (n < 0 || (n > 0 && !src) || !dst) return -1;
    std::vector<uint8_t> out;
    lz4_compress(static_cast<const uint8_t*>(src), n, out);
    if (static_cast<int64_t>(out.size()) > cap) return -1;
    std::memcpy(dst, out.data(), out.size());
    return static_cast<int64_t>(out.size());
}
SPHB_API int64_t sphb_lz4_decompress(const void* src, int64_t n, void* dst, int64_t cap) {
    if (n < 0 || !src || (cap > 0 && !dst)) return -1;
    return lz4_decompress(static_cast<const uint8_t*>(src), n, static_cast<uint8_t*>(dst), cap);
}

SPHB_API int sphb_cache_write_vec(const char* path, const void* data, int64_t count, int elem_size) {
    if (!path || count < 0 || elem_size <= 0 || (count > 0 && !data)) return fail(SPHB_ERR_INVALID, "cache_write_vec: bad arguments");
    return write_vec(path, static_cast<const uint8_t*>(data), count, elem_size);
}
SPHB_API int sphb_cache_read_vec(const char* path, void* out, int64_t cap, int elem_size, int64_t* count) {
    if (!path || elem_size <= 0 || !count) return fail(SPHB_ERR_INVALID, "cache_read_vec: bad arguments");
    return read_vec(path, static_cast<uint8_t*>(out), cap, elem_size, count);
}

SPHB_API int sphb_cache_write_graph(const char* base, int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, int symmetric) {
    if (!base || n < 0 || (!rowptr && k <= 0) || (n > 0 && (!ids || !dist))) return fail(SPHB_ERR_INVALID, "cache_write_graph: bad arguments");
    std::vector<int64_t> nns(static_cast<size_t>(n));
    for (int64_t r = 0; r < n; ++r) nns[r] = rowptr ? rowptr[r + 1] - rowptr[r] : k;
    const int64_t nnz = rowptr ? rowptr[n] : n * k;
    const std::string b(base);
    if (int rc = write_vec(b + "Distances.cache", reinterpret_cast<const uint8_t*>(dist), nnz, 4)) return rc;
    if (int rc = write_vec(b + "Indices.cache", reinterpret_cast<const uint8_t*>(ids), nnz, 8)) return rc;
    if (int rc = write_raw(b + "NNs.cache", nns.data(), nns.size() * 8)) return rc;
    const uint8_t sym = symmetric ? 1 : 0;
    return write_raw(b + "Symmetric.cache", &sym, 1);
}
SPHB_API int sphb_cache_read_graph(const char* base, int64_t cap_n, int64_t cap_nnz, int64_t* n, int64_t* nnz, int64_t* nns, int64_t* ids, float* dist,
                                   int* symmetric) {
    if (!base || !n || !nnz) return fail(SPHB_ERR_INVALID, "cache_read_graph: bad arguments");
    const std::string b(base);
    std::vector<uint8_t> raw;
    if (int rc = read_raw(b + "NNs.cache", raw)) return rc;
    *n = static_cast<int64_t>(raw.size() / 8);
    int64_t count = 0;
    if (int rc = read_vec(b + "Indices.cache", nullptr, 0, 8, &count)) return rc;
    *nnz = count;
    if (!nns && !ids && !dist) return SPHB_OK;
    if (!nns || !ids || !dist || cap_n < *n || cap_nnz < *nnz) return fail(SPHB_ERR_INVALID, "cache_read_graph: output capacity too small");
    std::memcpy(nns, raw.data(), static_cast<size_t>(*n) * 8);
    int64_t sum = 0;
    for (int64_t r = 0; r < *n; ++r) sum += nns[r];
    if (sum != *nnz) return fail(SPHB_ERR_INVALID, "cache_read_graph: neighbour counts do not add up to the index count");
    if (int rc = read_vec(b + "Indices.cache", reinterpret_cast<uint8_t*>(ids), cap_nnz, 8, &count)) return rc;
    if (int rc = read_vec(b + "Distances.cache", reinterpret_cast<uint8_t*>(dist), cap_nnz, 4, &count)) return rc;
    if (count != *nnz) return fail(SPHB_ERR_INVALID, "cache_read_graph: distances and indices differ in length");
    if (int rc = read_raw(b + "Symmetric.cache", raw)) return rc;
    if (symmetric) *symmetric = (!raw.empty() && raw[0]) ? 1 : 0;
    return SPHB_OK;
}

SPHB_API int sphb_cache_write_prob(const char* path, int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val) {
    if (!path || n < 0 || !rowptr || rowptr[0] != 0 || (rowptr[n] > 0 && (!col || !val))) return fail(SPHB_ERR_INVALID, "cache_write_prob: bad arguments");
    for (int64_t r = 0; r < n; ++r)
        if (rowptr[r + 1] < rowptr[r]) return fail(SPHB_ERR_INVALID, "cache_write_prob: rowptr not monotone");
    File fo(path, "wb");
    if (!fo.f) return fail(SPHB_ERR_INVALID, std::string("cache: cannot open ") + path + " for writing");
    const uint64_t total = 8 + static_cast<uint64_t>(n) * 8 + static_cast<uint64_t>(rowptr[n]) * 8, rows = static_cast<uint64_t>(n);
    std::vector<uint8_t> ser, buf;
    bool ok;
    if (total <= static_cast<uint64_t>(g_chunk_limit)) {  // calculateTotalBytesSparseMatHDI: fits while totalBytes <= INT_MAX (FileIO.cpp:111)
        const uint64_t chunks = 1;
        serialize_rows(rowptr, col, val, 0, n, ser);
        lz4_compress(ser.data(), static_cast<int64_t>(ser.size()), buf);
        const int32_t csize = static_cast<int32_t>(buf.size());
        ok = fo.put(&total, 8) && fo.put(&rows, 8) && fo.put(&chunks, 8) && fo.put(&csize, 4) && fo.put(buf.data(), buf.size());
    } else {
        const uint64_t chunks = static_cast<uint64_t>(num_chunks_for(total));
        ok = fo.put(&total, 8) && fo.put(&rows, 8) && fo.put(&chunks, 8);
        const int64_t base = n / static_cast<int64_t>(chunks), rem = n % static_cast<int64_t>(chunks);
        int64_t at = 0;
        for (uint64_t c = 0; ok && c < chunks; ++c) {
            const int64_t len = base + (static_cast<int64_t>(c) < rem ? 1 : 0);
            serialize_rows(rowptr, col, val, at, at + len, ser);
            lz4_compress(ser.data(), static_cast<int64_t>(ser.size()), buf);
            const uint32_t csize = static_cast<uint32_t>(buf.size()), osize = static_cast<uint32_t>(ser.size());
            ok = fo.put(&csize, 4) && fo.put(&osize, 4) && fo.put(buf.data(), buf.size());
            at += len;
        }
    }
    return ok ? SPHB_OK : fail(SPHB_ERR_INVALID, std::string("cache: short write to ") + path);
}
// rowptr == NULL: only *n and *nnz are filled (the blob is decompressed once to count the entries)
SPHB_API int sphb_cache_read_prob(const char* path, int64_t cap_n, int64_t cap_nnz, int64_t* n, int64_t* nnz, int64_t* rowptr, uint32_t* col, float* val) {
    if (!path || !n || !nnz || (rowptr && (!col || !val))) return fail(SPHB_ERR_INVALID, "cache_read_prob: bad arguments");
    File fi(path, "rb");
    if (!fi.f) return fail(SPHB_ERR_INVALID, std::string("cache: cannot open ") + path);
    uint64_t total = 0, rows = 0, chunks = 0;
    if (!fi.get(&total, 8) || !fi.get(&rows, 8) || !fi.get(&chunks, 8)) return fail(SPHB_ERR_INVALID, std::string("cache: ") + path + " has no header");
    if (rowptr) {
        if (cap_n < static_cast<int64_t>(rows)) return fail(SPHB_ERR_INVALID, "cache_read_prob: output capacity too small");
        rowptr[0] = 0;
    }
    int64_t row = 0, cnt = 0;
    std::vector<uint8_t> buf, ser;
    for (uint64_t c = 0; c < chunks; ++c) {
        uint32_t csize = 0, osize = static_cast<uint32_t>(total);
        if (!fi.get(&csize, 4)) return fail(SPHB_ERR_INVALID, std::string("cache: ") + path + " is truncated");
        if (chunks != 1 && !fi.get(&osize, 4)) return fail(SPHB_ERR_INVALID, std::string("cache: ") + path + " is truncated");
        buf.resize(csize);
        ser.resize(osize);
        if (!fi.get(buf.data(), buf.size())) return fail(SPHB_ERR_INVALID, std::string("cache: ") + path + " is truncated");
        if (lz4_decompress(buf.data(), csize, ser.data(), osize) != static_cast<int64_t>(osize))
            return fail(SPHB_ERR_INVALID, std::string("cache: ") + path + " does not decompress to its stated size");
        if (!parse_rows(ser.data(), osize, &row, &cnt, cap_n, cap_nnz, rowptr, col, val))
            return fail(SPHB_ERR_INVALID, std::string("cache: ") + path + " is malformed or larger than the output capacity");
    }
    if (row != static_cast<int64_t>(rows)) return fail(SPHB_ERR_INVALID, std::string("cache: ") + path + " holds another number of rows than its header says");
    *n = row;
    *nnz = cnt;
    return SPHB_OK;
}
}
