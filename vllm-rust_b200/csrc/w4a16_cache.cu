// w4a16_cache.cu -- on-disk cache of the resident (W4T) layout: w4a16_save_tiled / w4a16_load_tiled.
//
// The reference redoes its load-time repack on every start (AwqMarlinLinear::load_weights, awq_marlin.rs:580-667; the CPU
// version of the qweight transpose "took 5+ minutes" for a 4B model, awq_marlin.rs:202-206).  Here the repacked handle can
// be written once and read back with plain copies: a 160-byte header (magic, ABI version, full and local geometry,
// format, shard, a caller-supplied fingerprint of the source checkpoint, payload sizes, checksum) followed by
// qw | meta | bias | perm exactly as they sit in device memory.  Every header field is validated against the handle it is
// loaded into; a mismatch names the field.  Parity: unpack() after a cache round trip equals unpack() after a fresh
// repack bit for bit (tests/test_cache_gpu.py).
#include "../../include/w4a16_b200.h"

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "w4a16_internal.cuh"

namespace w4 {

struct TiledHeader {
    char magic[8];            // "W4TCACHE"
    uint32_t abi_version;     // W4A16_ABI_VERSION of the writer
    uint32_t header_bytes;    // sizeof(TiledHeader)
    int64_t K_full, N_full;   // checkpoint geometry
    int64_t K_local, N_local; // this shard
    int64_t k_off, n_off;
    int32_t group_size_cfg, format, has_bias;
    int32_t tp_mode, tp_rank, tp_size;
    int32_t has_perm, perm_external;
    uint64_t fingerprint;     // caller's identity of the source checkpoint (0 = unchecked)
    uint64_t qw_bytes, meta_bytes, bias_bytes, perm_bytes;
    uint64_t checksum;        // of the payload
    int32_t epilogue, reserved0;
    uint64_t reserved1;
};
static_assert(sizeof(TiledHeader) == 160, "W4T cache header layout");

// 64-bit multiply-xor hash over 8-byte words (FNV-1a style on words; the tail is zero-padded).
static uint64_t checksum64(const uint8_t *p, size_t n, uint64_t h) {
    const uint64_t prime = 0x100000001b3ull;
    size_t i = 0;
    for (; i + 8 <= n; i += 8) {
        uint64_t w;
        memcpy(&w, p + i, 8);
        h = (h ^ w) * prime;
    }
    if (i < n) {
        uint64_t w = 0;
        memcpy(&w, p + i, n - i);
        h = (h ^ w) * prime;
    }
    return h;
}

static TiledHeader header_of(const Linear &L, uint64_t fingerprint) {
    TiledHeader h{};
    memcpy(h.magic, "W4TCACHE", 8);
    h.abi_version = W4A16_ABI_VERSION;
    h.header_bytes = sizeof(TiledHeader);
    h.K_full = L.K_full; h.N_full = L.N_full; h.K_local = L.d.K; h.N_local = L.d.N;
    h.k_off = L.k_off; h.n_off = L.n_off;
    h.group_size_cfg = L.group_size_cfg; h.format = L.format; h.has_bias = L.has_bias ? 1 : 0;
    h.tp_mode = L.tp_mode; h.tp_rank = L.tp_rank; h.tp_size = L.tp_size;
    h.has_perm = L.perm ? 1 : 0; h.perm_external = L.perm_external ? 1 : 0;
    h.fingerprint = fingerprint;
    h.epilogue = L.epilogue;
    h.qw_bytes = (uint64_t)qw_bytes(L.d); h.meta_bytes = (uint64_t)meta_bytes(L.d);
    h.bias_bytes = L.has_bias ? (uint64_t)L.d.N * 2 : 0;
    h.perm_bytes = L.perm ? (uint64_t)(L.d.K + L.d.G) * 4 : 0;
    return h;
}

}  // namespace w4

using namespace w4;

extern "C" {

int w4a16_save_tiled(w4a16_handle_t hnd, const char *path, uint64_t fingerprint) {
    Linear *L = reinterpret_cast<Linear *>(hnd);
    if (!L || !L->loaded) return fail("save_tiled: weights not loaded");
    if (!path) return fail("save_tiled: null path");
    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != L->device) cudaSetDevice(L->device);
    TiledHeader h = header_of(*L, fingerprint);
    const size_t total = (size_t)(h.qw_bytes + h.meta_bytes + h.bias_bytes + h.perm_bytes);
    std::vector<uint8_t> buf(total);
    size_t off = 0;
    cudaError_t e = cudaDeviceSynchronize();   // the repack may still be in flight on some stream
    auto pull = [&](const void *src, size_t n) {
        if (n && e == cudaSuccess) e = cudaMemcpy(buf.data() + off, src, n, cudaMemcpyDeviceToHost);
        off += n;
    };
    pull(L->qw, (size_t)h.qw_bytes);
    pull(L->meta, (size_t)h.meta_bytes);
    pull(L->bias, (size_t)h.bias_bytes);
    pull(L->perm, (size_t)h.perm_bytes);
    if (prev >= 0 && prev != L->device) cudaSetDevice(prev);
    if (e != cudaSuccess) return fail(std::string("save_tiled: device read failed: ") + cudaGetErrorString(e));
    h.checksum = checksum64(buf.data(), total, 0xcbf29ce484222325ull);
    const std::string tmp = std::string(path) + ".tmp";
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f) return fail("save_tiled: cannot open '" + tmp + "' for writing");
    const bool ok = fwrite(&h, sizeof(h), 1, f) == 1 && (total == 0 || fwrite(buf.data(), 1, total, f) == total);
    const bool closed = fclose(f) == 0;
    if (!ok || !closed) { remove(tmp.c_str()); return fail("save_tiled: short write to '" + tmp + "'"); }
    if (rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return fail(std::string("save_tiled: cannot rename to '") + path + "'"); }
    return 0;
}

int w4a16_load_tiled(w4a16_handle_t hnd, const char *path, uint64_t fingerprint, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(hnd);
    if (!L) return fail("load_tiled: null handle");
    if (!path) return fail("load_tiled: null path");
    FILE *f = fopen(path, "rb");
    if (!f) return fail(std::string("load_tiled: cannot open '") + path + "'");
    TiledHeader h{};
    if (fread(&h, sizeof(h), 1, f) != 1) { fclose(f); return fail("load_tiled: file shorter than the header"); }
    auto bad = [&](const std::string &what) { fclose(f); return fail("load_tiled: " + what + " ('" + path + "')"); };
    if (memcmp(h.magic, "W4TCACHE", 8) != 0) return bad("not a W4T cache file (bad magic)");
    if (h.abi_version != W4A16_ABI_VERSION) return bad("written by ABI version " + std::to_string(h.abi_version) + ", this library is " + std::to_string(W4A16_ABI_VERSION));
    if (h.header_bytes != sizeof(TiledHeader)) return bad("header size mismatch");
    const TiledHeader want = header_of(*L, fingerprint);
#define W4_SAME(field)                                                                                              \
    if (h.field != want.field)                                                                                      \
        return bad(std::string("cached ") + #field + " = " + std::to_string((long long)h.field) + " but the handle has " + \
                   std::to_string((long long)want.field))
    W4_SAME(K_full); W4_SAME(N_full); W4_SAME(K_local); W4_SAME(N_local); W4_SAME(k_off); W4_SAME(n_off);
    W4_SAME(group_size_cfg); W4_SAME(format); W4_SAME(has_bias); W4_SAME(tp_mode); W4_SAME(tp_rank); W4_SAME(tp_size); W4_SAME(epilogue);
    W4_SAME(qw_bytes); W4_SAME(meta_bytes); W4_SAME(bias_bytes);
#undef W4_SAME
    if (fingerprint != 0 && h.fingerprint != fingerprint) return bad("the cache was built from a different checkpoint (fingerprint mismatch)");
    const uint64_t perm_want = h.has_perm ? (uint64_t)(L->d.K + L->d.G) * 4 : 0;
    if (h.perm_bytes != perm_want) return bad("permutation size mismatch");
    const size_t total = (size_t)(h.qw_bytes + h.meta_bytes + h.bias_bytes + h.perm_bytes);
    std::vector<uint8_t> buf(total);
    const size_t got = total ? fread(buf.data(), 1, total, f) : 0;
    uint8_t extra;
    const bool trailing = fread(&extra, 1, 1, f) == 1;
    fclose(f);
    if (got != total || trailing) return fail(std::string("load_tiled: payload size does not match the header ('") + path + "')");
    if (checksum64(buf.data(), total, 0xcbf29ce484222325ull) != h.checksum)
        return fail(std::string("load_tiled: checksum mismatch, the file is corrupt ('") + path + "')");

    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != L->device) cudaSetDevice(L->device);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    cudaError_t e = cudaSuccess;
    size_t off = 0;
    auto push = [&](void *dst, size_t n) {
        if (n && e == cudaSuccess) e = cudaMemcpyAsync(dst, buf.data() + off, n, cudaMemcpyHostToDevice, s);
        off += n;
    };
    push(L->qw, (size_t)h.qw_bytes);
    push(L->meta, (size_t)h.meta_bytes);
    push(L->bias, (size_t)h.bias_bytes);
    if (h.perm_bytes) {
        cudaFree(L->perm); L->perm = nullptr;
        if (e == cudaSuccess) e = cudaMalloc(&L->perm, (size_t)h.perm_bytes);
        push(L->perm, (size_t)h.perm_bytes);
    } else {
        cudaFree(L->perm); L->perm = nullptr;
    }
    L->perm_external = h.perm_external != 0;
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);   // `buf` is a local
    if (prev >= 0 && prev != L->device) cudaSetDevice(prev);
    if (e != cudaSuccess) return fail(std::string("load_tiled: device copy failed: ") + cudaGetErrorString(e));
    L->loaded = true;
    return 0;
}

}  // extern "C"
