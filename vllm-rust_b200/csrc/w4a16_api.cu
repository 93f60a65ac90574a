// w4a16_api.cu -- the extern "C" surface of libw4a16_b200.so (include/w4a16_b200.h).
// Handle lifetime, shape gates, load-time repack orchestration, forward dispatch,
// the row-parallel all-reduce and the parity hooks.  No compute lives here.
#include "../../include/w4a16_b200.h"

#include <dlfcn.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "w4a16_internal.cuh"
#include "w4a16_gemv_common.cuh"

namespace w4 {

static thread_local std::string t_last_error;
std::atomic<uint64_t> g_kernel_launches{0};

void set_error(const std::string &msg) { t_last_error = msg; }
int fail(const std::string &msg) {
    t_last_error = msg;
    return 1;
}

static std::mutex g_dev_mu;
static DeviceScratch g_dev[64];
static bool g_dev_init[64] = {};

int get_device_scratch(int device, DeviceScratch **out) {
    if (device < 0 || device >= 64) return fail("device index out of range");
    std::lock_guard<std::mutex> lk(g_dev_mu);
    if (!g_dev_init[device]) {
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0)
            return fail(std::string("no CUDA device available (libw4a16_b200 has no CPU fallback): ") +
                        cudaGetErrorString(e));
        if (device >= count) return fail("device index >= device count");
        cudaDeviceProp prop;
        W4_CUDA_OK(cudaGetDeviceProperties(&prop, device));
        if (prop.major != 10)
            return fail("libw4a16_b200 is built for sm_100a only; device " + std::to_string(device) + " is sm_" +
                        std::to_string(prop.major) + std::to_string(prop.minor));
        g_dev[device].sm_count = prop.multiProcessorCount;
        g_dev[device].cc_major = prop.major;
        g_dev[device].cc_minor = prop.minor;
        g_dev_init[device] = true;
    }
    *out = &g_dev[device];
    return 0;
}

static std::string g_dispatch = "auto";
constexpr int64_t DECODE_MAX_M = 32;

// marlin.rs:48, 308-341
static int check_shape(int64_t N, int64_t K, int gs) {
    if (gs != -1 && gs != 32 && gs != 64 && gs != 128)
        return fail("Marlin requires group_size in [-1, 32, 64, 128], got " + std::to_string(gs));
    if (N <= 0 || K <= 0) return fail("in_features and out_features must be positive");
    if (N % 64 != 0) return fail("out_features (" + std::to_string(N) + ") must be divisible by 64 for Marlin");
    if (K % 128 != 0) return fail("in_features (" + std::to_string(K) + ") must be divisible by 128 for Marlin");
    if (gs > 0 && K % gs != 0)
        return fail("in_features (" + std::to_string(K) + ") must be divisible by group_size (" + std::to_string(gs) + ")");
    return 0;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

static void free_linear(Linear *L) {
    if (!L) return;
    DeviceGuard g(L->device);
    cudaFree(L->qw); cudaFree(L->meta); cudaFree(L->bias); cudaFree(L->perm); cudaFree(L->tickets);
    cudaFree(L->splitk_scratch); cudaFree(L->x_gather); cudaFree(L->stage_x); cudaFree(L->stage_y);
    cudaFree(L->rowpar_f32); cudaFree(L->gv_tab);
    for (void *q : L->retired) cudaFree(q);
    delete L;
}

// One column part of a fused handle (w4a16_load_*_part): destination columns [n_begin, n_begin + n_part) from a
// checkpoint of width n_part.  whole(): the ordinary single-checkpoint load.
struct Part {
    int64_t n_begin = 0, n_part = 0;
    bool is_part = false;
};

static int finish_load(Linear *L, const void *bias, int bias_dtype, cudaStream_t stream, const Part &pt = Part()) {
    const bool swiglu = L->epilogue == W4A16_EPILOGUE_SWIGLU;
    if (L->has_bias) {
        if (!bias) return fail("load_weights: layer was created with bias but none was given");
        // bias is per output feature: a COLUMN shard keeps its slice, a ROW shard keeps all of it
        size_t esz = bias_dtype == W4A16_DTYPE_F32 ? 4 : 2;
        if (!pt.is_part) {
            const char *src = reinterpret_cast<const char *>(bias) + (size_t)L->n_off * esz;
            if (launch_convert_bias(src, bias_dtype, L->bias, L->d.N, stream)) return 1;
        } else if (swiglu) {
            if (launch_convert_bias(bias, bias_dtype, L->bias, pt.n_part, stream, pt.n_begin ? 1 : 0)) return 1;
        } else {
            if (launch_convert_bias(bias, bias_dtype, L->bias + pt.n_begin, pt.n_part, stream)) return 1;
        }
    }
    if (!pt.is_part) { L->loaded = true; return 0; }
    // a part: the handle is loaded once the parts cover every column exactly once
    for (int i = 0; i < L->n_parts_loaded; ++i)
        if (L->part_cols[i][0] == pt.n_begin && L->part_cols[i][1] == pt.n_part) return 0;   // reloaded in place
    L->part_cols[L->n_parts_loaded][0] = pt.n_begin; L->part_cols[L->n_parts_loaded][1] = pt.n_part;
    ++L->n_parts_loaded;
    int64_t covered = 0;
    for (int i = 0; i < L->n_parts_loaded; ++i) covered += L->part_cols[i][1];
    L->loaded = covered == L->d.N;
    return 0;
}

static int check_part(Linear *L, int64_t n_begin, int64_t n_part, Part *pt) {
    if (L->tp_mode != W4A16_TP_NONE) return fail("load_part: fused sibling parts need an unsharded handle");
    if (n_begin < 0 || n_part <= 0 || n_begin + n_part > L->d.N) return fail("load_part: columns out of range");
    if (L->epilogue == W4A16_EPILOGUE_SWIGLU) {
        if (n_part * 2 != L->d.N || (n_begin != 0 && n_begin != n_part))
            return fail("load_part: a SwiGLU handle takes exactly two parts of out_features / 2 columns: gate at 0, up at out_features / 2");
    } else if (n_begin % CHUNK_N || n_part % CHUNK_N) {
        return fail("load_part: part boundaries must be multiples of 16 columns");
    }
    bool again = false;   // checked BEFORE the repack writes anything
    for (int i = 0; i < L->n_parts_loaded; ++i) {
        if (L->part_cols[i][0] == n_begin && L->part_cols[i][1] == n_part) { again = true; continue; }
        if (n_begin < L->part_cols[i][0] + L->part_cols[i][1] && L->part_cols[i][0] < n_begin + n_part)
            return fail("load_part: the part overlaps columns that another part has filled");
    }
    if (!again && L->n_parts_loaded >= 8) return fail("load_part: more than 8 parts");
    pt->n_begin = n_begin; pt->n_part = n_part; pt->is_part = true;
    return 0;
}

// ---- NCCL, dlopen'ed (nccl.rs:451-547) ---------------------------------------------
struct UniqueId { char internal[W4A16_NCCL_UNIQUE_ID_BYTES]; };
struct Nccl {
    void *lib = nullptr;
    int (*GetUniqueId)(void *) = nullptr;
    int (*CommInitRank)(void **, int, UniqueId, int) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
static Nccl g_nccl;
static std::once_flag g_nccl_once;
static std::string g_nccl_err;

static int nccl_load() {
    std::call_once(g_nccl_once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (g_nccl.lib) break;
        }
        if (!g_nccl.lib) { g_nccl_err = std::string("dlopen(libnccl.so.2) failed: ") + dlerror(); return; }
        auto sym = [&](const char *s) { return dlsym(g_nccl.lib, s); };
        g_nccl.GetUniqueId = reinterpret_cast<int (*)(void *)>(sym("ncclGetUniqueId"));
        g_nccl.CommInitRank = reinterpret_cast<int (*)(void **, int, UniqueId, int)>(sym("ncclCommInitRank"));
        g_nccl.CommDestroy = reinterpret_cast<int (*)(void *)>(sym("ncclCommDestroy"));
        g_nccl.AllReduce = reinterpret_cast<int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t)>(sym("ncclAllReduce"));
        g_nccl.GetErrorString = reinterpret_cast<const char *(*)(int)>(sym("ncclGetErrorString"));
        if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllReduce)
            g_nccl_err = "libnccl.so.2 is missing a required symbol";
    });
    if (!g_nccl_err.empty()) return fail(g_nccl_err);
    return 0;
}
static int nccl_check(int rc, const char *what) {
    if (rc == 0) return 0;
    return fail(std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "NCCL error") +
                " (" + std::to_string(rc) + ")");
}
constexpr int NCCL_BF16 = 9, NCCL_SUM = 0;

// Scratch that scales with M is grown on demand but never freed before the handle is: a CUDA graph captured earlier may
// hold the old address, and a call inside a capture must not allocate at all (w4a16_reserve sizes it beforehand).
static int grow_guard(cudaStream_t stream, const char *what) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone)
        return fail(std::string(what) + ": scratch too small inside a stream capture -- call w4a16_reserve(handle, max_tokens) before capturing");
    return 0;
}
static int ensure_stage(Linear *L, int64_t M, cudaStream_t stream) {
    if (M <= L->stage_cap) return 0;
    if (grow_guard(stream, "forward_host")) return 1;
    if (L->stage_x) L->retired.push_back(L->stage_x);
    if (L->stage_y) L->retired.push_back(L->stage_y);
    L->stage_x = L->stage_y = nullptr; L->stage_cap = 0;
    W4_CUDA_OK(cudaMalloc(&L->stage_x, (size_t)M * L->d.K * 2));
    W4_CUDA_OK(cudaMalloc(&L->stage_y, (size_t)M * L->d.N * 2));
    L->stage_cap = M;
    return 0;
}
static int ensure_gather(Linear *L, int64_t M, cudaStream_t stream) {
    if (!L->perm || L->perm_external || M <= L->x_gather_cap) return 0;
    if (grow_guard(stream, "forward (act-order gather)")) return 1;
    if (L->x_gather) L->retired.push_back(L->x_gather);
    L->x_gather = nullptr; L->x_gather_cap = 0;
    W4_CUDA_OK(cudaMalloc(&L->x_gather, (size_t)M * L->d.K * 2));
    L->x_gather_cap = M;
    return 0;
}

// Which kernel family takes M tokens.  auto: M <= 10 -> the decode kernels (HBM-bound GEMV / register-ring mma.sync GEMM),
// 11..48 -> the mid-M kernel (mma.sync, every warp a row block x k slice, w4a16_midm.cu), above -> the prefill kernel
// (tcgen05; split-K fills the machine at few tokens).  Measured on the Qwen3-8B shapes:
// at M = 32 the prefill kernel takes 19 / 15 / 47 / 26 us against 23 / 28 / 74 / 48 us for the decode kernel, at
// M = 16 the decode kernel still wins (13 / 12 / 35 / 23 against 19 / 15 / 46 / 25 us).
enum { ROUTE_DECODE = 0, ROUTE_MIDM = 1, ROUTE_PREFILL = 2 };
static int env_int_api(const char *name, int dflt) { const char *v = getenv(name); return v ? atoi(v) : dflt; }
static int pick_route(const Linear &L, int64_t M) {
    if (g_dispatch == "prefill") return ROUTE_PREFILL;
    if (g_dispatch == "decode") return ROUTE_DECODE;
    static const int decode_max = std::min<int>((int)DECODE_MAX_M, exp_env_int("W4A16_DECODE_MAX_M", 16));
    if (M > 4 && midm_usable(L, M)) return ROUTE_MIDM;
    return M <= decode_max ? ROUTE_DECODE : ROUTE_PREFILL;
}

static inline int64_t out_width(const Linear &L) { return L.epilogue == W4A16_EPILOGUE_SWIGLU ? L.d.N / 2 : L.d.N; }

static int forward_local(Linear *L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                         cudaStream_t stream) {
    if (!L->loaded) return fail("forward: weights not loaded (call load_weights first)");
    if (M < 1) return fail("forward: M must be >= 1");
    if (!x || !y) return fail("forward: null x or y");
    if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(y) & 15))
        return fail("forward: x and y must be 16-byte aligned");
    if (L->perm && !L->perm_external) {  // act-order: tiled row i multiplies x[:, perm[i]]
        if (ensure_gather(L, M, stream)) return 1;
        if (launch_gather_x(x, L->perm, L->x_gather, M, L->d.K, stream)) return 1;
        x = L->x_gather;
    }
    int route = pick_route(*L, M);
    if (route == ROUTE_MIDM) {
        const int rc = launch_midm(*L, x, y, M, add_bias, stream);
        if (rc != 2) return rc;
        // 2: its split-K scratch for this stream would have to be allocated inside a stream capture -> the kernels that
        // took this M before it existed (a forward on the stream before capturing, or w4a16_reserve, avoids this)
        route = M <= 16 ? ROUTE_DECODE : ROUTE_PREFILL;
    }
    if (route == ROUTE_PREFILL) return launch_prefill(*L, x, y, M, add_bias, stream);
    for (int64_t m0 = 0; m0 < M; m0 += DECODE_MAX_M) {
        int64_t mm = std::min<int64_t>(DECODE_MAX_M, M - m0);
        if (launch_decode(*L, x + m0 * L->d.K, y + m0 * out_width(*L), mm, add_bias, stream)) return 1;
    }
    return 0;
}

// per-linear forward for callers outside this file (the chain's fallback, w4a16_chain.cu)
int forward_linear(Linear *L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, cudaStream_t stream) {
    const bool add_bias = L->tp_mode != W4A16_TP_ROW || L->tp_size == 1;
    return forward_local(L, x, y, M, add_bias, stream);
}

}  // namespace w4

using namespace w4;

extern "C" {

const char *w4a16_last_error(void) { return t_last_error.c_str(); }
int w4a16_abi_version(void) { return W4A16_ABI_VERSION; }
uint64_t w4a16_kernel_launches(void) { return g_kernel_launches.load(); }

int w4a16_check_shape(int64_t out_features, int64_t in_features, int group_size) {
    return check_shape(out_features, in_features, group_size);
}

int w4a16_create_shard(w4a16_handle_t *out, int64_t in_features, int64_t out_features, int group_size,
                       int format, int has_bias, int device, int tp_mode, int tp_rank, int tp_size) {
    if (!out) return fail("create: null out pointer");
    *out = nullptr;
    if (format < W4A16_FORMAT_AWQ || format > W4A16_FORMAT_GPTQ_SYM) return fail("create: unknown format");
    if (tp_mode < W4A16_TP_NONE || tp_mode > W4A16_TP_ROW) return fail("create: unknown tp_mode");
    if (tp_mode == W4A16_TP_NONE) { tp_rank = 0; tp_size = 1; }
    if (tp_size < 1 || tp_rank < 0 || tp_rank >= tp_size) return fail("create: bad tp_rank / tp_size");
    if (check_shape(out_features, in_features, group_size)) return 1;
    int64_t K = in_features, N = out_features, k_off = 0, n_off = 0;
    if (tp_mode == W4A16_TP_COLUMN) {
        if (N % tp_size) return fail("create: out_features not divisible by tp_size");
        N /= tp_size; n_off = tp_rank * N;
    } else if (tp_mode == W4A16_TP_ROW) {
        if (K % tp_size) return fail("create: in_features not divisible by tp_size");
        K /= tp_size; k_off = tp_rank * K;
        if (group_size > 0 && K % group_size) return fail("create: row shard is not a whole number of groups");
    }
    if (check_shape(N, K, group_size)) {  // the shard itself must pass the gate (SURVEY 8d gate check)
        set_error(std::string("create: shard fails the shape gate: ") + w4a16_last_error());
        return 1;
    }
    DeviceScratch *ds = nullptr;
    if (get_device_scratch(device, &ds)) return 1;
    DeviceGuard guard(device);
    Linear *L = new Linear();
    L->K_full = in_features; L->N_full = out_features; L->group_size_cfg = group_size;
    L->format = format; L->has_bias = has_bias != 0; L->device = device;
    L->tp_mode = tp_mode; L->tp_rank = tp_rank; L->tp_size = tp_size;
    L->k_off = k_off; L->n_off = n_off; L->scratch = ds;
    L->d.K = K; L->d.N = N; L->d.gs = group_size > 0 ? group_size : (int)K;
    L->d.KB = K / CHUNK_K; L->d.NR = N / CHUNK_N; L->d.G = K / L->d.gs;
    auto bail = [&](const std::string &m) { free_linear(L); return fail(m); };
#define W4_ALLOC(ptr, bytes)                                                                     \
    do {                                                                                         \
        cudaError_t _e = cudaMalloc(reinterpret_cast<void **>(&(ptr)), (bytes));                 \
        if (_e != cudaSuccess) return bail(std::string("cudaMalloc: ") + cudaGetErrorString(_e)); \
    } while (0)
    W4_ALLOC(L->qw, (size_t)qw_bytes(L->d));
    W4_ALLOC(L->meta, (size_t)meta_bytes(L->d));
    if (L->has_bias) W4_ALLOC(L->bias, (size_t)N * 2);
    int64_t sf = 0, nt = 0;
    decode_scratch_need(*L, ds->sm_count, &sf, &nt);
    W4_ALLOC(L->tickets, (size_t)nt * 4);
    if (cudaMemset(L->tickets, 0, (size_t)nt * 4) != cudaSuccess) return bail("cudaMemset(tickets) failed");
    if (sf > 0) { W4_ALLOC(L->splitk_scratch, (size_t)sf * 4); L->splitk_floats = sf; }
#undef W4_ALLOC
    if (gemv_build_table(*L, ds->sm_count)) { free_linear(L); return 1; }
    *out = reinterpret_cast<w4a16_handle_t>(L);
    return 0;
}

int w4a16_create(w4a16_handle_t *out, int64_t in_features, int64_t out_features, int group_size, int format,
                 int has_bias, int device) {
    return w4a16_create_shard(out, in_features, out_features, group_size, format, has_bias, device,
                              W4A16_TP_NONE, 0, 1);
}

int w4a16_destroy(w4a16_handle_t h) {
    free_linear(reinterpret_cast<Linear *>(h));
    return 0;
}

static int load_common(Linear *L, int format_expected_min, int format_expected_max, const uint32_t *qweight,
                       const uint32_t *qzeros, const void *scales, int scales_dtype, const int32_t *perm_dev,
                       const int32_t *grp_dev, cudaStream_t stream, const Part &pt = Part()) {
    if (L->format < format_expected_min || L->format > format_expected_max)
        return fail("load_weights: handle was created for a different checkpoint format");
    if (!qweight) return fail("load_weights: missing 'qweight'");
    if (!scales) return fail("load_weights: missing 'scales'");
    if (L->format != W4A16_FORMAT_GPTQ_SYM && !qzeros) return fail("load_weights: missing 'qzeros'");
    if (scales_dtype < W4A16_DTYPE_F16 || scales_dtype > W4A16_DTYPE_F32) return fail("load_weights: bad scales dtype");
    RepackSrc src;
    src.qweight = qweight; src.qzeros = qzeros; src.scales = scales; src.scales_dtype = scales_dtype;
    src.format = L->format; src.K_full = L->K_full; src.N_full = L->N_full;
    src.gs = L->group_size_cfg > 0 ? L->group_size_cfg : (int)L->K_full;
    src.G_full = L->K_full / src.gs;
    src.k_off = L->k_off; src.n_off = L->n_off;
    src.perm = perm_dev; src.row_group = grp_dev;
    if (pt.is_part) {   // the part's checkpoint is n_part wide and starts at its own column 0
        src.N_full = pt.n_part; src.n_off = 0;
        if (L->epilogue == W4A16_EPILOGUE_SWIGLU) { src.half = pt.n_begin ? 1 : 0; src.dst_R0 = 0; src.NR_part = L->d.NR; }
        else { src.dst_R0 = pt.n_begin / CHUNK_N; src.NR_part = pt.n_part / CHUNK_N; }
    } else if (L->epilogue != W4A16_EPILOGUE_NONE) {
        return fail("load_weights: a SwiGLU handle is loaded through w4a16_load_*_part (gate, then up)");
    }
    return launch_repack(*L, src, stream);
}

static int load_awq_impl(Linear *L, const Part &pt, const uint32_t *qweight, const uint32_t *qzeros, const void *scales,
                         int scales_dtype, const void *bias, int bias_dtype, cudaStream_t s) {
    DeviceGuard g(L->device);
    if (load_common(L, W4A16_FORMAT_AWQ, W4A16_FORMAT_AWQ, qweight, qzeros, scales, scales_dtype, nullptr, nullptr, s, pt))
        return 1;
    return finish_load(L, bias, bias_dtype, s, pt);
}

int w4a16_load_awq(w4a16_handle_t h, const uint32_t *qweight, const uint32_t *qzeros, const void *scales,
                   int scales_dtype, const void *bias, int bias_dtype, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("load_awq: null handle");
    return load_awq_impl(L, Part(), qweight, qzeros, scales, scales_dtype, bias, bias_dtype, reinterpret_cast<cudaStream_t>(stream));
}

int w4a16_load_awq_part(w4a16_handle_t h, int64_t n_begin, int64_t n_part, const uint32_t *qweight, const uint32_t *qzeros,
                        const void *scales, int scales_dtype, const void *bias, int bias_dtype, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("load_awq_part: null handle");
    Part pt;
    if (check_part(L, n_begin, n_part, &pt)) return 1;
    return load_awq_impl(L, pt, qweight, qzeros, scales, scales_dtype, bias, bias_dtype, reinterpret_cast<cudaStream_t>(stream));
}

int w4a16_set_epilogue(w4a16_handle_t h, int epilogue) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("set_epilogue: null handle");
    if (epilogue != W4A16_EPILOGUE_NONE && epilogue != W4A16_EPILOGUE_SWIGLU) return fail("set_epilogue: unknown epilogue");
    if (L->loaded || L->n_parts_loaded) return fail("set_epilogue: must be called before any weights are loaded");
    if (epilogue == W4A16_EPILOGUE_SWIGLU && L->tp_mode != W4A16_TP_NONE) return fail("set_epilogue: a SwiGLU handle must be unsharded");
    L->epilogue = epilogue;
    return 0;
}

static int load_gptq_impl(Linear *L, const Part &pt, const uint32_t *qweight, const uint32_t *qzeros, const void *scales,
                          int scales_dtype, const int32_t *g_idx, const void *bias, int bias_dtype, cudaStream_t s) {
    DeviceGuard g(L->device);
    const int32_t *perm_dev = nullptr, *grp_dev = nullptr;
    if (pt.is_part && L->n_parts_loaded > 0 && (g_idx != nullptr) != !L->part_g_idx.empty())
        return fail("load_gptq_part: every part of a fused handle must bring the same g_idx (one has it, another does not)");
    if (g_idx) {
        // g_idx is the FULL [K_full] tensor.  The rows are sorted by group over the whole K (stable), THEN sharded
        // ("permute at load before sharding"): a ROW shard keeps sorted rows [k_off, k_off + K_local), i.e.
        // whole groups [k_off / gs, ...), and expects its x columns in that order (perm_external: no gather at
        // forward; w4a16_row_perm hands the permutation to the caller, who folds it into the producing layer).
        const int64_t Kf = L->K_full, K = L->d.K;
        const int gs = L->d.gs;
        const int64_t Gf = Kf / gs, G = L->d.G;
        std::vector<int32_t> gi((size_t)Kf);
        W4_CUDA_OK(cudaMemcpyAsync(gi.data(), g_idx, (size_t)Kf * 4, cudaMemcpyDeviceToHost, s));
        W4_CUDA_OK(cudaStreamSynchronize(s));
        bool trivial = true;
        for (int64_t k = 0; k < Kf; ++k) {
            if (gi[k] < 0 || gi[k] >= Gf) return fail("load_gptq: g_idx entry out of range");
            if (gi[k] != k / gs) trivial = false;
        }
        if (pt.is_part) {   // sibling projections share their input: one permutation of x serves all parts
            if (L->n_parts_loaded > 0 && L->part_g_idx != gi)
                return fail("load_gptq_part: every part of a fused handle must bring the same g_idx");
            L->part_g_idx = gi;
        }
        if (!trivial && pt.is_part && L->perm) {
            perm_dev = L->perm; grp_dev = L->perm + L->d.K;   // built by the first part
        } else if (!trivial) {
            std::vector<int32_t> full((size_t)Kf);
            std::vector<int64_t> count((size_t)Gf, 0);
            for (int64_t k = 0; k < Kf; ++k) ++count[gi[k]];
            for (int64_t gg = 0; gg < Gf; ++gg)
                if (count[gg] != gs) return fail("load_gptq: g_idx must assign exactly group_size rows to every group");
            std::vector<int64_t> next((size_t)Gf);
            for (int64_t gg = 0; gg < Gf; ++gg) next[gg] = gg * gs;
            for (int64_t k = 0; k < Kf; ++k) full[next[gi[k]]++] = (int32_t)k;   // sorted row i <- checkpoint row full[i]
            // the shard's slice, stored relative to k_off (the repack adds k_off back; a ROW shard's entries may be
            // negative or beyond K_local: they are checkpoint rows, not local x columns)
            std::vector<int32_t> perm((size_t)K + (size_t)G);
            for (int64_t i = 0; i < K; ++i) perm[i] = (int32_t)(full[L->k_off + i] - L->k_off);
            for (int64_t gg = 0; gg < G; ++gg) perm[K + gg] = (int32_t)(L->k_off / gs + gg);  // sorted block -> checkpoint group
            cudaFree(L->perm); L->perm = nullptr;
            W4_CUDA_OK(cudaMalloc(&L->perm, perm.size() * 4));
            W4_CUDA_OK(cudaMemcpyAsync(L->perm, perm.data(), perm.size() * 4, cudaMemcpyHostToDevice, s));
            W4_CUDA_OK(cudaStreamSynchronize(s));  // perm is a local; the copy must finish before it dies
            perm_dev = L->perm; grp_dev = L->perm + K;
            L->perm_external = L->tp_mode == W4A16_TP_ROW && L->tp_size > 1;
        }
    }
    if (load_common(L, W4A16_FORMAT_GPTQ_ZP, W4A16_FORMAT_GPTQ_SYM, qweight, qzeros, scales, scales_dtype, perm_dev,
                    grp_dev, s, pt))
        return 1;
    return finish_load(L, bias, bias_dtype, s, pt);
}

int w4a16_load_gptq(w4a16_handle_t h, const uint32_t *qweight, const uint32_t *qzeros, const void *scales,
                    int scales_dtype, const int32_t *g_idx, const void *bias, int bias_dtype,
                    w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("load_gptq: null handle");
    return load_gptq_impl(L, Part(), qweight, qzeros, scales, scales_dtype, g_idx, bias, bias_dtype, reinterpret_cast<cudaStream_t>(stream));
}

int w4a16_load_gptq_part(w4a16_handle_t h, int64_t n_begin, int64_t n_part, const uint32_t *qweight, const uint32_t *qzeros,
                         const void *scales, int scales_dtype, const int32_t *g_idx, const void *bias, int bias_dtype,
                         w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("load_gptq_part: null handle");
    Part pt;
    if (check_part(L, n_begin, n_part, &pt)) return 1;
    return load_gptq_impl(L, pt, qweight, qzeros, scales, scales_dtype, g_idx, bias, bias_dtype, reinterpret_cast<cudaStream_t>(stream));
}

int w4a16_forward_into(w4a16_handle_t h, const void *x_bf16, void *y_bf16, int64_t M, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("forward: null handle");
    DeviceGuard g(L->device);
    // a ROW shard returns its partial sum without bias (bias is added after the reduce)
    bool add_bias = L->tp_mode != W4A16_TP_ROW || L->tp_size == 1;
    return forward_local(L, reinterpret_cast<const __nv_bfloat16 *>(x_bf16), reinterpret_cast<__nv_bfloat16 *>(y_bf16),
                         M, add_bias, reinterpret_cast<cudaStream_t>(stream));
}

int w4a16_forward_host(w4a16_handle_t h, const void *x_host, void *y_host, int64_t M, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("forward_host: null handle");
    if (M < 1) return fail("forward_host: M must be >= 1");
    DeviceGuard g(L->device);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (ensure_stage(L, M, s)) return 1;
    W4_CUDA_OK(cudaMemcpyAsync(L->stage_x, x_host, (size_t)M * L->d.K * 2, cudaMemcpyHostToDevice, s));
    if (w4a16_forward_into(h, L->stage_x, L->stage_y, M, stream)) return 1;
    W4_CUDA_OK(cudaMemcpyAsync(y_host, L->stage_y, (size_t)M * out_width(*L) * 2, cudaMemcpyDeviceToHost, s));
    W4_CUDA_OK(cudaStreamSynchronize(s));
    return 0;
}

int w4a16_forward_row_allreduce(w4a16_handle_t h, const void *x_bf16, void *y_bf16, int64_t M, w4a16_comm_t comm,
                                w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("forward_row_allreduce: null handle");
    DeviceGuard g(L->device);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const bool single = L->tp_size == 1 || L->tp_mode != W4A16_TP_ROW;
    if (single)  // world-size-1 short-circuit (nccl.rs:1339-1341)
        return forward_local(L, reinterpret_cast<const __nv_bfloat16 *>(x_bf16),
                             reinterpret_cast<__nv_bfloat16 *>(y_bf16), M, true, s);
    if (!comm) return fail("forward_row_allreduce: null communicator for tp_size > 1");
    if (nccl_load()) return 1;
    if (forward_local(L, reinterpret_cast<const __nv_bfloat16 *>(x_bf16), reinterpret_cast<__nv_bfloat16 *>(y_bf16), M,
                      false, s))
        return 1;
    if (nccl_check(g_nccl.AllReduce(y_bf16, y_bf16, (size_t)M * L->d.N, NCCL_BF16, NCCL_SUM, comm, s), "ncclAllReduce"))
        return 1;
    if (L->has_bias) return launch_add_bias(reinterpret_cast<__nv_bfloat16 *>(y_bf16), L->bias, M, L->d.N, s);
    return 0;
}

// ---- symmetric peer buffers + fused GEMV / all-reduce ------------------------------------------
// Replaces the NCCL call of RowParallelLinear::forward (parallel_layers.rs:295-304, nccl.rs:1338-1364) and the
// dormant kernels/custom_allreduce.cu for the decode sizes: the GEMV's epilogue stores its fp32 partials into
// every peer's buffer over NVLink and sums the world's partials itself (w4a16_gemv_imma.cu).
__global__ void peer_fill_kernel(uint32_t *p, size_t n, uint32_t v) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

struct PeerCtx {
    int rank = 0, world = 1, device = 0;
    uint32_t cap = 0;            // fp32 words per (buffer, source)
    size_t bytes = 0;
    float *local = nullptr;
    float *base[PEER_MAX_WORLD] = {};
    bool opened[PEER_MAX_WORLD] = {};
    bool attached = false;
    uint32_t *epoch_dev = nullptr;   // [0] launch counter, [1] finished-CTA ticket (device, this rank only)
};

// The device-side view of an attached peer group, for callers outside this file (the chain kernel's all-reduce).
extern "C++" {
namespace w4 {
int peer_make_view(void *peer, PeerView *v) {
    PeerCtx *c = reinterpret_cast<PeerCtx *>(peer);
    if (!c) return fail("peer group: null");
    if (!c->attached) return fail("peer group: buffers not attached");
    *v = PeerView{};
    for (int r = 0; r < c->world; ++r) v->base[r] = c->base[r];
    v->world = c->world; v->rank = c->rank; v->cap = c->cap;
    v->flags = 0u;
    v->epoch_ptr = c->epoch_dev; v->ticket = reinterpret_cast<int *>(c->epoch_dev + 1);
    return 0;
}
}  // namespace w4
}  // extern "C++"

int w4a16_peer_create(w4a16_peer_t *out, int rank, int world_size, int64_t max_partial_floats) {
    if (!out) return fail("peer_create: null out");
    if (world_size < 2 || world_size > PEER_MAX_WORLD) return fail("peer_create: world_size must be in 2..8");
    if (rank < 0 || rank >= world_size) return fail("peer_create: bad rank");
    if (max_partial_floats < 64 || max_partial_floats > ((int64_t)1 << 26)) return fail("peer_create: bad capacity");
    PeerCtx *c = new PeerCtx();
    c->rank = rank; c->world = world_size;
    W4_CUDA_OK(cudaGetDevice(&c->device));
    c->cap = (uint32_t)((max_partial_floats + 63) & ~(int64_t)63);
    c->bytes = (size_t)3 * world_size * c->cap * sizeof(float);
    if (cudaMalloc(&c->local, c->bytes) != cudaSuccess) { delete c; return fail("peer_create: cudaMalloc failed"); }
    {   // every word starts as the "empty" sentinel
        const size_t words = c->bytes / 4;
        peer_fill_kernel<<<(unsigned)((words + 255) / 256), 256>>>(reinterpret_cast<uint32_t *>(c->local), words, PEER_EMPTY);
        W4_CUDA_OK(cudaGetLastError());
    }
    W4_CUDA_OK(cudaMalloc(&c->epoch_dev, 2 * sizeof(uint32_t)));
    W4_CUDA_OK(cudaMemset(c->epoch_dev, 0, 2 * sizeof(uint32_t)));
    W4_CUDA_OK(cudaDeviceSynchronize());
    c->base[rank] = c->local;
    *out = reinterpret_cast<w4a16_peer_t>(c);
    return 0;
}

int w4a16_peer_export(w4a16_peer_t peer, void *handle_out) {
    PeerCtx *c = reinterpret_cast<PeerCtx *>(peer);
    if (!c || !handle_out) return fail("peer_export: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == W4A16_PEER_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t hnd;
    W4_CUDA_OK(cudaIpcGetMemHandle(&hnd, c->local));
    memcpy(handle_out, &hnd, sizeof(hnd));
    return 0;
}

int w4a16_peer_attach(w4a16_peer_t peer, const void *all_handles) {
    PeerCtx *c = reinterpret_cast<PeerCtx *>(peer);
    if (!c || !all_handles) return fail("peer_attach: null argument");
    if (c->attached) return fail("peer_attach: already attached");
    DeviceGuard g(c->device);
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        cudaIpcMemHandle_t hnd;
        memcpy(&hnd, reinterpret_cast<const char *>(all_handles) + (size_t)r * sizeof(hnd), sizeof(hnd));
        void *ptr = nullptr;
        W4_CUDA_OK(cudaIpcOpenMemHandle(&ptr, hnd, cudaIpcMemLazyEnablePeerAccess));
        c->base[r] = reinterpret_cast<float *>(ptr);
        c->opened[r] = true;
    }
    c->attached = true;
    return 0;
}

int w4a16_peer_destroy(w4a16_peer_t peer) {
    PeerCtx *c = reinterpret_cast<PeerCtx *>(peer);
    if (!c) return 0;
    DeviceGuard g(c->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < c->world; ++r)
        if (c->opened[r]) cudaIpcCloseMemHandle(c->base[r]);
    cudaFree(c->local);
    cudaFree(c->epoch_dev);
    delete c;
    return 0;
}

int w4a16_forward_row_fused(w4a16_handle_t h, const void *x_bf16, void *y_bf16, int64_t M, w4a16_peer_t peer,
                            w4a16_comm_t comm, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    PeerCtx *c = reinterpret_cast<PeerCtx *>(peer);
    if (!L) return fail("forward_row_fused: null handle");
    if (L->tp_size == 1 || L->tp_mode != W4A16_TP_ROW || !c)
        return w4a16_forward_row_allreduce(h, x_bf16, y_bf16, M, comm, stream);
    if (!c->attached) return fail("forward_row_fused: peer buffers not attached");
    if (c->world != L->tp_size || c->rank != L->tp_rank) return fail("forward_row_fused: peer group does not match the shard");
    DeviceGuard g(L->device);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    PeerView v{};
    for (int r = 0; r < c->world; ++r) v.base[r] = c->base[r];
    v.world = c->world; v.rank = c->rank; v.cap = c->cap;
    v.flags = exp_env_int("W4A16_FUSED_EARLY_PDL", 0) ? 1u : 0u;
    v.epoch_ptr = c->epoch_dev; v.ticket = reinterpret_cast<int *>(c->epoch_dev + 1);
    // Every rank takes the same branch: the test depends on shape, M and the group only.  The launch counter lives
    // in device memory, so the fused kernel can be captured into a CUDA graph and replayed.
    if (M > PEER_TOK || !L->loaded || !gemv_imma_peer_usable(*L, M, v) || env_int_api("W4A16_NO_FUSED_AR", 0))
        return w4a16_forward_row_allreduce(h, x_bf16, y_bf16, M, comm, stream);
    if (!x_bf16 || !y_bf16) return fail("forward_row_fused: null x or y");
    const __nv_bfloat16 *x = reinterpret_cast<const __nv_bfloat16 *>(x_bf16);
    if (L->perm && !L->perm_external) {
        if (ensure_gather(L, M, s)) return 1;
        if (launch_gather_x(x, L->perm, L->x_gather, M, L->d.K, s)) return 1;
        x = L->x_gather;
    }
    return launch_gemv_imma_peer(*L, x, reinterpret_cast<__nv_bfloat16 *>(y_bf16), M, true, s, &v);
}

int w4a16_nccl_unique_id(void *unique_id_out) {
    if (!unique_id_out) return fail("nccl_unique_id: null buffer");
    if (nccl_load()) return 1;
    return nccl_check(g_nccl.GetUniqueId(unique_id_out), "ncclGetUniqueId");
}
int w4a16_nccl_comm_init(w4a16_comm_t *comm_out, int world_size, const void *unique_id, int rank) {
    if (!comm_out || !unique_id) return fail("nccl_comm_init: null argument");
    if (nccl_load()) return 1;
    UniqueId id;
    memcpy(&id, unique_id, sizeof(id));
    void *c = nullptr;
    if (nccl_check(g_nccl.CommInitRank(&c, world_size, id, rank), "ncclCommInitRank")) return 1;
    *comm_out = c;
    return 0;
}
int w4a16_nccl_comm_destroy(w4a16_comm_t comm) {
    if (!comm) return 0;
    if (nccl_load()) return 1;
    return nccl_check(g_nccl.CommDestroy(comm), "ncclCommDestroy");
}

int w4a16_reserve(w4a16_handle_t h, int64_t max_tokens) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L) return fail("reserve: null handle");
    if (max_tokens < 1) return fail("reserve: max_tokens must be >= 1");
    DeviceGuard g(L->device);
    if (ensure_gather(L, max_tokens, nullptr)) return 1;
    return ensure_stage(L, max_tokens, nullptr);
}

int w4a16_dequant_into(w4a16_handle_t h, void *w_bf16, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L || !L->loaded) return fail("dequant_into: weights not loaded");
    DeviceGuard g(L->device);
    return launch_dequant(*L, reinterpret_cast<__nv_bfloat16 *>(w_bf16), reinterpret_cast<cudaStream_t>(stream));
}
int w4a16_dequant_d1_into(w4a16_handle_t h, const void *scales, int scales_dtype, void *w_f16, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L || !L->loaded) return fail("dequant_d1_into: weights not loaded");
    if (!scales || !w_f16) return fail("dequant_d1_into: null scales or output");
    if (scales_dtype < W4A16_DTYPE_F16 || scales_dtype > W4A16_DTYPE_F32) return fail("dequant_d1_into: bad scales dtype");
    if (L->n_parts_loaded) return fail("dequant_d1_into: a fused handle has one scales tensor per part; use the per-sibling handles");
    DeviceGuard g(L->device);
    return launch_dequant_d1(*L, scales, scales_dtype, reinterpret_cast<__half *>(w_f16), reinterpret_cast<cudaStream_t>(stream));
}
int w4a16_row_perm(w4a16_handle_t h, int32_t *perm_out_host) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L || !L->loaded) return fail("row_perm: weights not loaded");
    if (!perm_out_host) return fail("row_perm: null output");
    const int64_t K = L->d.K;
    if (!L->perm) {
        for (int64_t i = 0; i < K; ++i) perm_out_host[i] = (int32_t)(L->k_off + i);
        return 0;
    }
    DeviceGuard g(L->device);
    W4_CUDA_OK(cudaMemcpy(perm_out_host, L->perm, (size_t)K * 4, cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < K; ++i) perm_out_host[i] += (int32_t)L->k_off;
    return 0;
}
int w4a16_unpack_into(w4a16_handle_t h, uint8_t *q, uint8_t *z, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L || !L->loaded) return fail("unpack_into: weights not loaded");
    DeviceGuard g(L->device);
    return launch_unpack(*L, q, z, reinterpret_cast<cudaStream_t>(stream));
}
int w4a16_scales_into(w4a16_handle_t h, void *scales_bf16, w4a16_stream_t stream) {
    Linear *L = reinterpret_cast<Linear *>(h);
    if (!L || !L->loaded) return fail("scales_into: weights not loaded");
    DeviceGuard g(L->device);
    return launch_scales_out(*L, reinterpret_cast<__nv_bfloat16 *>(scales_bf16), reinterpret_cast<cudaStream_t>(stream));
}

int64_t w4a16_in_features(w4a16_handle_t h) { return h ? reinterpret_cast<Linear *>(h)->d.K : -1; }
int64_t w4a16_out_features(w4a16_handle_t h) {   /* a SwiGLU handle writes silu(gate) * up: half as wide as its weights */
    if (!h) return -1;
    const Linear *L = reinterpret_cast<Linear *>(h);
    return L->epilogue == W4A16_EPILOGUE_SWIGLU ? L->d.N / 2 : L->d.N;
}
int64_t w4a16_weight_rows(w4a16_handle_t h) { return h ? reinterpret_cast<Linear *>(h)->d.N : -1; }
int w4a16_has_bias(w4a16_handle_t h) { return h ? (int)reinterpret_cast<Linear *>(h)->has_bias : 0; }
int64_t w4a16_weight_bytes(w4a16_handle_t h) {
    if (!h) return -1;
    Linear *L = reinterpret_cast<Linear *>(h);
    return qw_bytes(L->d) + meta_bytes(L->d);
}
const char *w4a16_kernel_for(w4a16_handle_t h, int64_t M) {
    if (!h) return "none";
    const int route = pick_route(*reinterpret_cast<Linear *>(h), M);
    if (route == ROUTE_MIDM) return "midm_mma";
    if (route == ROUTE_PREFILL) return "prefill_umma";
    return decode_kernel_name(*reinterpret_cast<Linear *>(h), std::min<int64_t>(M, DECODE_MAX_M));
}
int w4a16_debug_stamps(uint64_t *out, int64_t words, uint64_t *launches) {
    if (!out) return fail("debug_stamps: null buffer");
    unsigned long long n = 0;
    if (gemv_debug_read(reinterpret_cast<unsigned long long *>(out), words, &n)) return 1;
    if (launches) *launches = n;
    return 0;
}
int w4a16_set_dispatch(const char *mode) {
    if (!mode) return fail("set_dispatch: null mode");
    std::string m(mode);
    if (m != "auto" && m != "decode" && m != "prefill") return fail("set_dispatch: unknown mode '" + m + "'");
    g_dispatch = m;
    return 0;
}

}  // extern "C"
