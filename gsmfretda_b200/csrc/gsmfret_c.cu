// gsmfret_c.cu — runtime behind include/gsmfret_c.h: the role of the reference's class mc
// (/root/reference/src/mc.cu:199-829) as a C ABI.  All device memory is owned here;
// nothing on the evaluation path allocates once the slot has seen its largest (N, P).
#include "../../include/gsmfret_c.h"
#include "pda_kernels.h"

#include <cuda_profiler_api.h>
#include <dlfcn.h>
#include <unistd.h>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace pda;

// ------------------------------------------------------------------------------------
// error plumbing
// ------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

// the same, for the library's other translation units (gsmfret_static.cu); not exported
__attribute__((visibility("hidden"))) int gsm_detail_fail(int code, const char* text) { return fail(code, "%s", text); }

#define CK(stmt)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (stmt);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return fail(GSM_ECUDA, "%s returned %s(%d) at %s:%d", #stmt, cudaGetErrorString(e_), \
                        (int)e_, __FILE__, __LINE__);                                         \
    } while (0)

// ------------------------------------------------------------------------------------
// NCCL through dlopen: only burst-split runs need it
// ------------------------------------------------------------------------------------
struct Id128 { char b[128]; };  // ncclUniqueId is passed by value: 128 bytes
struct NcclApi {
    void* h = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, Id128, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;

static int nccl_load()
{
    if (g_nccl.h) return GSM_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(GSM_ENCCL, "dlopen(libnccl.so.2) failed: %s", dlerror());
    g_nccl.GetUniqueId = (int (*)(void*))dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank = (int (*)(void**, int, Id128, int))dlsym(h, "ncclCommInitRank");
    g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(h, "ncclAllReduce");
    g_nccl.CommDestroy = (int (*)(void*))dlsym(h, "ncclCommDestroy");
    g_nccl.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy)
        return fail(GSM_ENCCL, "libnccl lacks a required symbol");
    g_nccl.h = h;
    return GSM_OK;
}
// cuRAND's host API hands out the Sobol tables mc::mc fetches (/root/reference/src/mc.cu:312-314);
// only the reference-draws mode needs them
static int load_sobol_tables(gsm_ctx* c);

constexpr int kNcclUint64 = 5;  // ncclUint64 in nccl.h's ncclDataType_t
constexpr int kNcclSum = 0;     // ncclSum

// ------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------
struct ResultHeader {
    unsigned long long flags;
    unsigned long long stamp[4];   // %globaltimer of the last block: start, posted, all seen, done
    unsigned long long reserved;
};

// bench.py declares these sizes as the bytes that cross the bus per evaluation (PARAM_RAW_BYTES, RESULT_HEADER_BYTES)
static_assert(sizeof(ParamRaw) == 616 && sizeof(ResultHeader) == 48, "update bench.py's declared copy sizes and DESIGN.md");

constexpr int kDirectResultMaxP = 16;   // up to this many vectors the kernel writes results to host memory itself

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int n = 0;           // s_n[sid]
    int P = 0;           // vectors staged in h_params
    int N = -1;          // bursts of the current range (setBurstBd)
    uint32_t epoch = 0;
    // parameter staging
    int cap_P = 0;
    ParamRaw* h_params = nullptr;  // pinned
    ParamRaw* d_params = nullptr;
    bool params_resident = false;
    // working set + results
    size_t cap_res = 0;  // bytes of one result block
    unsigned long long* d_whist = nullptr;
    unsigned long long* d_counters = nullptr;
    unsigned char* d_res = nullptr;
    unsigned char* h_res = nullptr;  // pinned mirror: header | hist[P][nbp] | chi2[P]
    unsigned char* h_res_dev = nullptr;  // the same memory as the device sees it (mapped)
    int res_P = 0, res_nbp = 0;
    // optional per-sample outputs
    size_t cap_samples = 0;
    float* d_mcE = nullptr; float* h_mcE = nullptr;
    long long* d_cnt = nullptr; long long* h_cnt = nullptr;
    int32_t* d_steps = nullptr; int32_t* h_steps = nullptr;
    bool ran = false;
    // results of this slot live in vector alias_p of slot alias_sid (it was a member of a coalesced
    // launch, gsm_run_coalesced) until the slot runs by itself again; -1: its own
    int alias_sid = -1, alias_p = 0;
    bool per_vector_epochs = false;   // the staged vectors carry their own draw epochs
    uint32_t peer_seq = 0;   // evaluations exchanged on this slot so far
    // reference-draws mode: the per-thread generators of the reference (devQStates / devStates)
    ulonglong2* d_sob_state = nullptr;
    uint4* d_xs_state = nullptr;
    long long sob_NN = 0;      // threads the states were set up for (the reference re-seeds when N changes)
    long long sob_cap = 0;
};

struct gsm_ctx {
    int devid = 0;
    unsigned flags = 0;
    char uuid[33] = {0};
    std::vector<Slot> slots;
    int R = 4;  // reSampleTimes default, /root/reference/src/mc.cu:308
    uint32_t seed = 0x9E3779B9u;
    int nb = 200, left_pad = 2;
    // burst data
    bool inited = false;
    int64_t B = 0, n_photons = 0, g_offset = 0;
    bool packed = false, with_bg = false;
    int layout = 0;  // 0 raw int64 ticks, 1 packed words, 2 packed words + index
    int mode = 0;    // GSM_MODE_REPLAY / GSM_MODE_NORMAL
    float clk_p = 0, bg_ad = 0, bg_dd = 0;
    BurstRec* d_bursts = nullptr;
    int32_t* d_pk = nullptr;
    int32_t* d_pm = nullptr;
    uint32_t* d_nx = nullptr;
    uint32_t* d_cw = nullptr;
    int64_t* d_ad = nullptr; int64_t* d_dd = nullptr;
    float* d_sg = nullptr;
    float* d_ohist = nullptr;
    std::vector<float> h_ohist;
    LaunchCfg cfg{0, 0, 0};
    int occ_force = -1;   // GSM_OCC=0/1 pins the register budget of the evaluation kernel (measurements); -1: by workload
    int chunk_force = 0, span_force = 0, tail_force = -1;   // GSM_CHUNK (log2), GSM_SPAN, GSM_TAIL: launch shape (tests)
    std::atomic<uint64_t> launches{0};
    void* comm = nullptr;
    int rank = 0, world = 1;
    // fused peer exchange (gsm_peer_export / gsm_peer_connect)
    unsigned long long* peer_base = nullptr;   // this rank's symmetric block, all slots
    size_t peer_bytes = 0;
    size_t peer_slot_words = 0;                // u64 words per slot
    int peer_cap_P = 0, peer_nbp = 0, peer_world = 0;
    bool peer_on = false;
    unsigned long long* peer_map[kMaxPeers] = {nullptr};
    bool peer_ipc[kMaxPeers] = {false};
    uint32_t peer_timeout_ms = 20000;
    // reference-draws mode: cuRAND's scrambled Sobol64 tables (20000 dimensions)
    unsigned long long* d_sob_dv = nullptr;
    unsigned long long* d_sob_sc = nullptr;
};

// what travels between ranks to map each other's symmetric block
struct PeerHandle {
    uint32_t magic;
    uint32_t pid;
    int32_t devid;
    int32_t world;
    uint64_t ptr;
    uint64_t bytes;
    cudaIpcMemHandle_t ipc;   // 64 bytes
    unsigned char uuid[16];   // the GPU, whatever index it has in the exporting process
    unsigned char pad[128 - 32 - 16 - sizeof(cudaIpcMemHandle_t)];
};
static_assert(sizeof(PeerHandle) == GSM_PEER_HANDLE_BYTES, "peer handle is 128 bytes");
constexpr uint32_t kPeerMagic = 0x47534d50u;  // "GSMP"
static void peer_release(gsm_ctx* c)
{
    for (int r = 0; r < kMaxPeers; ++r) {
        if (c->peer_ipc[r] && c->peer_map[r]) cudaIpcCloseMemHandle(c->peer_map[r]);
        c->peer_map[r] = nullptr;
        c->peer_ipc[r] = false;
    }
    if (c->peer_base) cudaFree(c->peer_base);
    c->peer_base = nullptr;
    c->peer_bytes = 0; c->peer_on = false; c->peer_world = 0; c->peer_cap_P = 0;
    cudaGetLastError();
}


static inline size_t res_bytes(int P, int nbp)
{
    size_t b = sizeof(ResultHeader) + (size_t)P * nbp * sizeof(unsigned long long) + (size_t)P * sizeof(float);
    return (b + 255) & ~(size_t)255;
}
static inline unsigned long long* res_hist(unsigned char* base) { return (unsigned long long*)(base + sizeof(ResultHeader)); }
static inline float* res_chi2(unsigned char* base, int P, int nbp)
{
    return (float*)(base + sizeof(ResultHeader) + (size_t)P * nbp * sizeof(unsigned long long));
}

// ------------------------------------------------------------------------------------
// host parameter math: the roles of genMatK / genP_i2j / genMatP
// (/root/reference/src/eigenhelper.cpp:4-54) and of the split in mc::set_params (mc.cu:792-799)
// ------------------------------------------------------------------------------------
static bool expand_params(int n, const float* args, int n_args, ParamRaw* out)
{
    if (n < 2 || n > kMaxStates) return false;  // genMatK refuses an empty rate list (n = 1)
    if (n_args != n * n + n) return false;
    memset(out, 0, sizeof(*out));
    out->n = n;
    const float* rates = args + n;
    for (int s = 0; s < n; ++s) {
        out->E[s] = args[s];
        out->v[s] = args[n * n + s];
    }
    // rates are listed row by row with the diagonal left out; K is stored column-major
    int next = 0;
    for (int row = 0; row < n; ++row)
        for (int col = 0; col < n; ++col)
            if (row != col) out->K[row + col * n] = rates[next++];
    for (int col = 0; col < n; ++col) {
        float diag = 0.0f;
        for (int row = 0; row < n; ++row)
            if (row != col) diag -= out->K[row + col * n];
        out->K[col + col * n] = diag;
    }
    for (int col = 0; col < n; ++col) {
        float leave = 0.0f;
        for (int row = 0; row < n; ++row)
            if (row != col) leave += out->K[row + col * n];
        for (int row = 0; row < n; ++row)
            out->P_i2j[row + col * n] = row != col ? out->K[row + col * n] / leave : 0.0f;
    }
    for (int row = 0; row < n; ++row) {
        float inflow = 0.0f;
        for (int col = 0; col < n; ++col) inflow += out->K[row + col * n];
        out->p[row] = inflow - out->K[row + row * n];
    }
    return true;
}

// ------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------
static int check_sid(const gsm_ctx* c, int sid)
{
    if (!c) return fail(GSM_EINVAL, "null context");
    if (sid < 0 || sid >= (int)c->slots.size()) return fail(GSM_EINVAL, "stream id %d out of range [0,%d)", sid, (int)c->slots.size());
    return GSM_OK;
}

static int ensure_params_cap(gsm_ctx* c, Slot& s, int P)
{
    if (P <= s.cap_P) return GSM_OK;
    CK(cudaStreamSynchronize(s.stream));
    if (s.h_params) CK(cudaFreeHost(s.h_params));
    if (s.d_params) CK(cudaFree(s.d_params));
    s.h_params = nullptr; s.d_params = nullptr; s.cap_P = 0;
    CK(cudaHostAlloc((void**)&s.h_params, (size_t)P * sizeof(ParamRaw), cudaHostAllocDefault));
    CK(cudaMalloc((void**)&s.d_params, (size_t)P * sizeof(ParamRaw)));
    s.cap_P = P;
    (void)c;
    return GSM_OK;
}

static int ensure_result_cap(gsm_ctx* c, Slot& s, int P, int64_t N)
{
    const int nbp = c->nb + 2;
    const size_t need = res_bytes(P, nbp);
    if (need > s.cap_res) {
        CK(cudaStreamSynchronize(s.stream));
        if (s.d_res) CK(cudaFree(s.d_res));
        if (s.h_res) CK(cudaFreeHost(s.h_res));
        if (s.d_whist) CK(cudaFree(s.d_whist));
        s.d_res = nullptr; s.h_res = nullptr; s.d_whist = nullptr; s.cap_res = 0;
        CK(cudaMalloc((void**)&s.d_res, need));
        CK(cudaHostAlloc((void**)&s.h_res, need, cudaHostAllocMapped));
        CK(cudaHostGetDevicePointer((void**)&s.h_res_dev, s.h_res, 0));
        CK(cudaMalloc((void**)&s.d_whist, need));
        CK(cudaMemsetAsync(s.d_whist, 0, need, s.stream));
        CK(cudaMemsetAsync(s.d_res, 0, need, s.stream));
        memset(s.h_res, 0, need);
        s.cap_res = need;
    }
    if (!s.d_counters) {
        CK(cudaMalloc((void**)&s.d_counters, 4 * sizeof(unsigned long long)));
        CK(cudaMemsetAsync(s.d_counters, 0, 4 * sizeof(unsigned long long), s.stream));
    }
    const size_t samples = (size_t)P * (size_t)N * (size_t)c->R;
    if ((c->flags & (GSM_KEEP_MCE | GSM_KEEP_COUNTS)) && samples > s.cap_samples) {
        CK(cudaStreamSynchronize(s.stream));
        if (c->flags & GSM_KEEP_MCE) {
            if (s.d_mcE) CK(cudaFree(s.d_mcE));
            if (s.h_mcE) CK(cudaFreeHost(s.h_mcE));
            CK(cudaMalloc((void**)&s.d_mcE, samples * sizeof(float)));
            CK(cudaHostAlloc((void**)&s.h_mcE, samples * sizeof(float), cudaHostAllocDefault));
        }
        if (c->flags & GSM_KEEP_COUNTS) {
            if (s.d_cnt) CK(cudaFree(s.d_cnt));
            if (s.h_cnt) CK(cudaFreeHost(s.h_cnt));
            if (s.d_steps) CK(cudaFree(s.d_steps));
            if (s.h_steps) CK(cudaFreeHost(s.h_steps));
            CK(cudaMalloc((void**)&s.d_cnt, samples * sizeof(long long)));
            CK(cudaHostAlloc((void**)&s.h_cnt, samples * sizeof(long long), cudaHostAllocDefault));
            CK(cudaMalloc((void**)&s.d_steps, samples * sizeof(int32_t)));
            CK(cudaHostAlloc((void**)&s.h_steps, samples * sizeof(int32_t), cudaHostAllocDefault));
        }
        s.cap_samples = samples;
    }
    return GSM_OK;
}

static int build_observed(gsm_ctx* c)
{
    const int nbp = c->nb + 2;
    unsigned long long* d_h = nullptr;
    CK(cudaMalloc((void**)&d_h, nbp * sizeof(unsigned long long)));
    CK(cudaMemset(d_h, 0, nbp * sizeof(unsigned long long)));
    if (c->d_ohist) CK(cudaFree(c->d_ohist));
    c->d_ohist = nullptr;
    CK(cudaMalloc((void**)&c->d_ohist, nbp * sizeof(float)));
    CK(launch_hist_f32(c->d_sg, c->B, c->nb, d_h, 0));
    CK(launch_u64_to_f32(d_h, c->d_ohist, nbp, 0));
    c->launches += (c->B ? 2 : 1);
    c->h_ohist.assign(nbp, 0.0f);
    CK(cudaMemcpy(c->h_ohist.data(), c->d_ohist, nbp * sizeof(float), cudaMemcpyDeviceToHost));
    CK(cudaFree(d_h));
    return GSM_OK;
}

static int load_sobol_tables(gsm_ctx* c)
{
    if (c->d_sob_dv) return GSM_OK;
    void* h = dlopen("libcurand.so.10", RTLD_NOW);
    if (!h) h = dlopen("libcurand.so", RTLD_NOW);
    if (!h) return fail(GSM_ESTATE, "reference-draws mode needs libcurand (dlopen: %s)", dlerror());
    typedef int (*GetDv)(unsigned long long (**)[64], int);
    typedef int (*GetSc)(unsigned long long**);
    GetDv get_dv = (GetDv)dlsym(h, "curandGetDirectionVectors64");
    GetSc get_sc = (GetSc)dlsym(h, "curandGetScrambleConstants64");
    if (!get_dv || !get_sc) return fail(GSM_ESTATE, "libcurand lacks the Sobol64 table getters");
    unsigned long long (*dv)[64] = nullptr;
    unsigned long long* sc = nullptr;
    const int kScrambledJoeKuo6_64 = 104;   // CURAND_SCRAMBLED_DIRECTION_VECTORS_64_JOEKUO6 (curand.h)
    if (get_dv(&dv, kScrambledJoeKuo6_64) != 0 || get_sc(&sc) != 0 || !dv || !sc)
        return fail(GSM_ESTATE, "curandGetDirectionVectors64 / curandGetScrambleConstants64 failed");
    CK(cudaSetDevice(c->devid));
    CK(cudaMalloc((void**)&c->d_sob_dv, 20000 * 64 * sizeof(unsigned long long)));
    CK(cudaMalloc((void**)&c->d_sob_sc, 20000 * sizeof(unsigned long long)));
    CK(cudaMemcpy(c->d_sob_dv, dv, 20000 * 64 * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->d_sob_sc, sc, 20000 * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    return GSM_OK;
}

// ------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------
extern "C" {

int gsm_abi_version(void) { return GSM_ABI_VERSION; }
const char* gsm_last_error(void) { return g_err; }

int gsm_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int d = 0; d < n; ++d) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10) ok++;
    }
    return ok;
}

int gsm_create(int devid, int stream_num, unsigned flags, gsm_ctx** out)
{
    if (!out) return fail(GSM_EINVAL, "out is null");
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(GSM_ENODEVICE, "no CUDA device: this library has no CPU path");
    }
    if (devid < 0 || devid >= ndev) return fail(GSM_EINVAL, "gpu id set error! devid %d of %d", devid, ndev);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, devid));
    if (prop.major != 10)
        return fail(GSM_ENODEVICE, "device %d is sm_%d%d; the kernels are built for sm_100a only", devid, prop.major, prop.minor);
    CK(cudaSetDevice(devid));
    if (stream_num < 0 || stream_num > 256) return fail(GSM_EINVAL, "stream_num %d", stream_num);
    if (stream_num == 0) stream_num = 2 * (prop.asyncEngineCount > 0 ? prop.asyncEngineCount : 1);
    gsm_ctx* c = new gsm_ctx();
    c->devid = devid;
    c->flags = flags;
    if (const char* occ = getenv("GSM_OCC")) {
        if (occ[0] == '0' || occ[0] == '1') c->occ_force = occ[0] - '0';
    }
    if (const char* e = getenv("GSM_CHUNK")) c->chunk_force = atoi(e);
    if (const char* e = getenv("GSM_SPAN")) c->span_force = atoi(e);
    if (const char* e = getenv("GSM_TAIL")) c->tail_force = atoi(e) != 0;
    for (int i = 0; i < 16; ++i) snprintf(c->uuid + 2 * i, 3, "%02x", (unsigned char)prop.uuid.bytes[i]);
    c->slots.resize(stream_num);
    for (int sid = 0; sid < stream_num; ++sid) {
        Slot& s = c->slots[sid];
        s.epoch = (uint32_t)sid << 24;
        cudaError_t e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreate(&s.ev0);
        if (e == cudaSuccess) e = cudaEventCreate(&s.ev1);
        if (e != cudaSuccess) {
            gsm_destroy(c);
            return fail(GSM_ECUDA, "stream/event creation returned %s(%d)", cudaGetErrorString(e), (int)e);
        }
    }
    *out = c;
    return GSM_OK;
}

void gsm_destroy(gsm_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->devid);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    for (Slot& s : c->slots)
        if (s.stream) cudaStreamSynchronize(s.stream);
    peer_release(c);
    for (Slot& s : c->slots) {
        cudaFreeHost(s.h_params); cudaFree(s.d_params);
        cudaFree(s.d_whist); cudaFree(s.d_counters); cudaFree(s.d_res); cudaFreeHost(s.h_res);
        cudaFree(s.d_sob_state); cudaFree(s.d_xs_state);
        cudaFree(s.d_mcE); cudaFreeHost(s.h_mcE);
        cudaFree(s.d_cnt); cudaFreeHost(s.h_cnt);
        cudaFree(s.d_steps); cudaFreeHost(s.h_steps);
        if (s.ev0) cudaEventDestroy(s.ev0);
        if (s.ev1) cudaEventDestroy(s.ev1);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    cudaFree(c->d_bursts); cudaFree(c->d_pk); cudaFree(c->d_pm); cudaFree(c->d_nx); cudaFree(c->d_cw);
    cudaFree(c->d_ad); cudaFree(c->d_dd);
    cudaFree(c->d_sg); cudaFree(c->d_ohist);
    cudaFree(c->d_sob_dv); cudaFree(c->d_sob_sc);
    cudaGetLastError();
    delete c;
}

int gsm_profiler(int on)
{
    CK(on ? cudaProfilerStart() : cudaProfilerStop());
    return GSM_OK;
}

int gsm_set_device(gsm_ctx* c)
{
    if (!c) return fail(GSM_EINVAL, "null context");
    CK(cudaSetDevice(c->devid));
    return GSM_OK;
}

int gsm_stream_count(const gsm_ctx* c) { return c ? (int)c->slots.size() : 0; }
void* gsm_stream(const gsm_ctx* c, int sid) { return (c && sid >= 0 && sid < (int)c->slots.size()) ? (void*)c->slots[sid].stream : nullptr; }
int gsm_device_id(const gsm_ctx* c) { return c ? c->devid : -1; }
int gsm_device_uuid(const gsm_ctx* c, char out[33])
{
    if (!c || !out) return fail(GSM_EINVAL, "null argument");
    memcpy(out, c->uuid, 33);
    return GSM_OK;
}

int gsm_set_resample_times(gsm_ctx* c, int R)
{
    if (!c || R < 1 || R > (1 << 20)) return fail(GSM_EINVAL, "reSampleTimes %d", R);
    c->R = R;
    return GSM_OK;
}
int gsm_resample_times(const gsm_ctx* c) { return c ? c->R : 0; }
int gsm_set_seed(gsm_ctx* c, uint32_t seed)
{
    if (!c) return fail(GSM_EINVAL, "null context");
    c->seed = seed;
    return GSM_OK;
}

int gsm_set_mode(gsm_ctx* c, int mode)
{
    if (!c) return fail(GSM_EINVAL, "null context");
    if (mode != GSM_MODE_REPLAY && mode != GSM_MODE_NORMAL && mode != GSM_MODE_REFERENCE)
        return fail(GSM_EINVAL, "unknown mode %d", mode);
    if (mode != c->mode) {
        for (Slot& s : c->slots) CK(cudaStreamSynchronize(s.stream));
        if (mode == GSM_MODE_REFERENCE) {
            int r = load_sobol_tables(c);
            if (r) return r;
            for (Slot& s : c->slots) s.sob_NN = 0;   // a fresh mc: setup_kernel runs before the first evaluation
        }
        c->mode = mode;
        if (c->inited) c->cfg = eval_launch_cfg(c->devid, c->layout, c->with_bg, c->nb, c->mode);
    }
    return GSM_OK;
}
int gsm_mode(const gsm_ctx* c) { return c ? c->mode : -1; }

int gsm_set_hist(gsm_ctx* c, int n_bins, int left_pad)
{
    if (!c || n_bins < 1 || n_bins + 2 > 1024 || left_pad < 0 || left_pad > n_bins)
        return fail(GSM_EINVAL, "histogram of %d bins, left pad %d", n_bins, left_pad);
    CK(cudaSetDevice(c->devid));
    const bool changed = n_bins != c->nb;
    c->nb = n_bins;
    c->left_pad = left_pad;
    if (changed) {
        for (Slot& s : c->slots) {  // result blocks are laid out per bin count
            CK(cudaStreamSynchronize(s.stream));
            if (s.d_whist) CK(cudaMemset(s.d_whist, 0, s.cap_res));
            s.ran = false;
        }
        if (c->inited) {
            c->cfg = eval_launch_cfg(c->devid, c->layout, c->with_bg, c->nb, c->mode);
            return build_observed(c);
        }
    }
    return GSM_OK;
}

int gsm_init_data(gsm_ctx* c, int64_t B, const int64_t* istart, const int64_t* start,
                  const int64_t* stop, const int32_t* ph_count, int64_t n_photons,
                  const int64_t* burst_ad, const int64_t* burst_dd, const float* T,
                  const float* SgDivSr, float clk_p, float bg_ad_rate, float bg_dd_rate,
                  int64_t g_offset)
{
    if (!c) return fail(GSM_EINVAL, "null context");
    if (B < 0 || n_photons < 0) return fail(GSM_EINVAL, "negative sizes");
    if (B > 0 && (!istart || !start || !stop || !ph_count || !T || !SgDivSr)) return fail(GSM_EINVAL, "null burst array");
    if (n_photons > 0 && (!burst_ad || !burst_dd)) return fail(GSM_EINVAL, "null photon array");
    if (!(clk_p > 0.0f) || !std::isfinite(clk_p)) return fail(GSM_EINVAL, "clk_p %g", (double)clk_p);
    if (!std::isfinite(bg_ad_rate) || !std::isfinite(bg_dd_rate) || bg_ad_rate < 0.0f || bg_dd_rate < 0.0f)
        return fail(GSM_EINVAL, "background rates %g %g", (double)bg_ad_rate, (double)bg_dd_rate);
    const bool with_bg = !((double)bg_dd_rate < 1e-4);  // the reference's switch, mc.cu:153
    std::vector<int64_t> poff((size_t)B);
    int64_t pk_total = 0;
    double max_lambda = 0.0;
    int64_t max_ph = 0;
    for (int64_t b = 0; b < B; ++b) {
        const int64_t ph = ph_count[b];
        if (ph < 0 || istart[b] < 0 || istart[b] + ph > n_photons)
            return fail(GSM_EINVAL, "burst %lld: photons [%lld, %lld) outside [0, %lld)", (long long)b,
                        (long long)istart[b], (long long)(istart[b] + ph), (long long)n_photons);
        poff[b] = pk_total;
        if (ph > max_ph) max_ph = ph;
        pk_total += (ph + 3) & ~(int64_t)3;
        if (with_bg) {
            double span = (double)(stop[b] - start[b]) * (double)clk_p;
            double lam = span * (double)(bg_ad_rate > bg_dd_rate ? bg_ad_rate : bg_dd_rate);
            if (lam > max_lambda) max_lambda = lam;
        }
    }
    if (pk_total >= (1ll << 34)) return fail(GSM_ERANGE, "%lld photon words: the packed offset covers 2^34", (long long)pk_total);
    if (with_bg && !(max_lambda < 900.0))
        return fail(GSM_ERANGE, "background mean %g per burst: the Poisson draw contract covers lambda < 900", max_lambda);
    CK(cudaSetDevice(c->devid));
    for (Slot& s : c->slots) CK(cudaStreamSynchronize(s.stream));
    cudaFree(c->d_bursts); cudaFree(c->d_pk); cudaFree(c->d_pm); cudaFree(c->d_nx); cudaFree(c->d_cw);
    cudaFree(c->d_ad); cudaFree(c->d_dd); cudaFree(c->d_sg);
    c->d_bursts = nullptr; c->d_pk = nullptr; c->d_pm = nullptr; c->d_nx = nullptr; c->d_cw = nullptr;
    c->d_ad = nullptr; c->d_dd = nullptr; c->d_sg = nullptr;
    c->inited = false;

    int64_t *d_start = nullptr, *d_stop = nullptr, *d_istart = nullptr, *d_poff = nullptr;
    int32_t* d_ph = nullptr;
    float* d_T = nullptr;
    unsigned int* d_bad = nullptr;
    const size_t Bn = (size_t)(B > 0 ? B : 1), Pn = (size_t)(n_photons > 0 ? n_photons : 1);
    auto free_tmp = [&]() { cudaFree(d_start); cudaFree(d_stop); cudaFree(d_istart); cudaFree(d_poff); cudaFree(d_ph); cudaFree(d_T); cudaFree(d_bad); };
#define CKT(stmt) do { cudaError_t e2_ = (stmt); if (e2_ != cudaSuccess) { free_tmp(); return fail(GSM_ECUDA, "%s returned %s(%d) at %s:%d", #stmt, cudaGetErrorString(e2_), (int)e2_, __FILE__, __LINE__); } } while (0)
    CKT(cudaMalloc((void**)&d_start, Bn * 8)); CKT(cudaMalloc((void**)&d_stop, Bn * 8));
    CKT(cudaMalloc((void**)&d_istart, Bn * 8)); CKT(cudaMalloc((void**)&d_poff, Bn * 8));
    CKT(cudaMalloc((void**)&d_ph, Bn * 4)); CKT(cudaMalloc((void**)&d_T, Bn * 4));
    CKT(cudaMalloc((void**)&d_bad, 4));
    CKT(cudaMalloc((void**)&c->d_sg, Bn * 4));
    CKT(cudaMalloc((void**)&c->d_bursts, Bn * sizeof(BurstRec)));
    CKT(cudaMalloc((void**)&c->d_ad, Pn * 8)); CKT(cudaMalloc((void**)&c->d_dd, Pn * 8));
    if (B > 0) {
        CKT(cudaMemcpy(d_start, start, B * 8, cudaMemcpyHostToDevice));
        CKT(cudaMemcpy(d_stop, stop, B * 8, cudaMemcpyHostToDevice));
        CKT(cudaMemcpy(d_istart, istart, B * 8, cudaMemcpyHostToDevice));
        CKT(cudaMemcpy(d_poff, poff.data(), B * 8, cudaMemcpyHostToDevice));
        CKT(cudaMemcpy(d_ph, ph_count, B * 4, cudaMemcpyHostToDevice));
        CKT(cudaMemcpy(d_T, T, B * 4, cudaMemcpyHostToDevice));
        CKT(cudaMemcpy(c->d_sg, SgDivSr, B * 4, cudaMemcpyHostToDevice));
    }
    if (n_photons > 0) {
        CKT(cudaMemcpy(c->d_ad, burst_ad, n_photons * 8, cudaMemcpyHostToDevice));
        CKT(cudaMemcpy(c->d_dd, burst_dd, n_photons * 8, cudaMemcpyHostToDevice));
    }
    CKT(cudaMemset(d_bad, 0, 4));
    bool packed = !(c->flags & GSM_FORCE_RAW64);
    // the next-photon distances are 16-bit: longer bursts keep the photon-by-photon scan
    bool indexed = packed && !(c->flags & GSM_NO_INDEX) && max_ph <= 65535;
    PackArgs pa{};
    pa.start = d_start; pa.stop = d_stop; pa.istart = d_istart; pa.ph = d_ph; pa.T = d_T;
    pa.ad = c->d_ad; pa.dd = c->d_dd; pa.poff = d_poff; pa.bursts = c->d_bursts;
    pa.bad = d_bad; pa.B = B;
    if (packed) {
        const size_t words = (size_t)(pk_total > 0 ? pk_total : 4);
        CKT(cudaMalloc((void**)&c->d_pk, words * 4));
        pa.pk = c->d_pk;
        if (indexed) {
            CKT(cudaMalloc((void**)&c->d_pm, words * 4));
            CKT(cudaMalloc((void**)&c->d_nx, words * 4));
            CKT(cudaMalloc((void**)&c->d_cw, words * 4));
            pa.pm = c->d_pm; pa.nx = c->d_nx; pa.cw = c->d_cw;
        }
        CKT(launch_pack(pa, 0));
        c->launches += (B ? (indexed ? 2 : 1) : 0);
        unsigned int bad = 0;
        CKT(cudaMemcpy(&bad, d_bad, 4, cudaMemcpyDeviceToHost));
        if (bad) {  // ticks too far from their burst start for 30 bits: keep the raw layout
            packed = false; indexed = false;
            cudaFree(c->d_pk); cudaFree(c->d_pm); cudaFree(c->d_nx); cudaFree(c->d_cw);
            c->d_pk = nullptr; c->d_pm = nullptr; c->d_nx = nullptr; c->d_cw = nullptr;
        }
    }
    if (!packed) {
        pa.pk = nullptr; pa.pm = nullptr; pa.nx = nullptr; pa.cw = nullptr;
        CKT(launch_pack(pa, 0));
        c->launches += (B ? 1 : 0);
    } else {
        cudaFree(c->d_ad); cudaFree(c->d_dd);
        c->d_ad = nullptr; c->d_dd = nullptr;
    }
    CKT(cudaDeviceSynchronize());
#undef CKT
    free_tmp();
    c->B = B; c->n_photons = n_photons; c->g_offset = g_offset;
    c->packed = packed; c->with_bg = with_bg;
    c->layout = packed ? (indexed ? 2 : 1) : 0;
    c->clk_p = clk_p; c->bg_ad = bg_ad_rate; c->bg_dd = bg_dd_rate;
    c->cfg = eval_launch_cfg(c->devid, c->layout, with_bg, c->nb, c->mode);
    c->inited = true;
    for (Slot& s : c->slots) { s.N = -1; s.ran = false; }
    return build_observed(c);
}

int64_t gsm_n_bursts(const gsm_ctx* c) { return c ? c->B : 0; }
int gsm_is_packed(const gsm_ctx* c) { return c && c->packed ? 1 : 0; }
int gsm_layout(const gsm_ctx* c) { return c ? c->layout : -1; }

int gsm_observed_hist(gsm_ctx* c, float* out, int n_bins)
{
    if (!c || !out) return fail(GSM_EINVAL, "null argument");
    if (!c->inited) return fail(GSM_ESTATE, "init_data has not run");
    if (n_bins != c->nb) return fail(GSM_EINVAL, "n_bins %d != %d", n_bins, c->nb);
    memcpy(out, c->h_ohist.data(), (size_t)n_bins * sizeof(float));
    return GSM_OK;
}

int gsm_set_observed_hist(gsm_ctx* c, const float* ohist, int n_bins)
{
    if (!c || !ohist) return fail(GSM_EINVAL, "null argument");
    if (!c->inited) return fail(GSM_ESTATE, "init_data has not run");
    if (n_bins != c->nb) return fail(GSM_EINVAL, "n_bins %d != %d", n_bins, c->nb);
    CK(cudaSetDevice(c->devid));
    for (Slot& s : c->slots) CK(cudaStreamSynchronize(s.stream));
    memcpy(c->h_ohist.data(), ohist, (size_t)n_bins * sizeof(float));
    CK(cudaMemcpy(c->d_ohist, ohist, (size_t)n_bins * sizeof(float), cudaMemcpyHostToDevice));
    return GSM_OK;
}

int gsm_get_max_conversion_capacity(gsm_ctx* c, int) { return c ? 1 : 0; }

int gsm_set_nstates(gsm_ctx* c, int n, int sid)
{
    int r = check_sid(c, sid);
    if (r) return r;
    if (n < 1 || n > kMaxStates) return fail(GSM_EINVAL, "n states %d outside [1,%d]", n, kMaxStates);
    Slot& s = c->slots[sid];
    const int prev = s.n;
    s.n = n;
    return prev;
}

int gsm_set_params(gsm_ctx* c, int n, int sid, const float* args, int n_args)
{
    int r = check_sid(c, sid);
    if (r) return r;
    if (!args) return fail(GSM_EINVAL, "null args");
    CK(cudaSetDevice(c->devid));
    Slot& s = c->slots[sid];
    r = ensure_params_cap(c, s, 1);
    if (r) return r;
    CK(cudaStreamSynchronize(s.stream));  // the staging buffer may still feed the previous run
    if (!expand_params(n, args, n_args, &s.h_params[0])) {
        // the slot holds no vector now: a run on it is refused instead of evaluating the previous one
        s.P = 0;
        s.ran = false;
        s.params_resident = false;
        fail(GSM_EINVAL, "set_params: %d values for %d states (need n*n+n, 2 <= n <= %d)", n_args, n, kMaxStates);
        return 0;
    }
    s.n = n;
    s.P = 1;
    s.params_resident = false;
    return 1;
}

int gsm_set_burst_bd(gsm_ctx* c, int cstart, int cstop, int sid)
{
    int r = check_sid(c, sid);
    if (r) return r;
    if (!c->inited) return fail(GSM_ESTATE, "init_data has not run");
    const int64_t szb = c->B;
    int64_t rstart = cstart, rstop = cstop;
    if (cstop >= szb || cstop < 0) rstop = szb;
    if (cstart >= rstop) rstart = 0;
    int64_t N = rstop - rstart;
    if (cstart == -1) N = -1;
    c->slots[sid].N = (int)N;
    return (int)N;
}

int gsm_init_randstate(gsm_ctx* c, int, int sid)
{
    int r = check_sid(c, sid);
    if (r) return r;
    // counter-based modes: nothing to build, the draw epoch already advances with every run.
    // reference-draws mode: mc::init_randstate rebuilds the Sobol and xor128 states from scratch
    if (c->mode == GSM_MODE_REFERENCE) c->slots[sid].sob_NN = 0;
    return GSM_OK;
}

int gsm_set_params_buff(gsm_ctx* c, int, int, int sid) { return check_sid(c, sid); }

int gsm_set_epoch(gsm_ctx* c, int sid, uint32_t epoch)
{
    int r = check_sid(c, sid);
    if (r) return r;
    c->slots[sid].epoch = epoch;
    return GSM_OK;
}

static int run_common(gsm_ctx* c, int N, int sid, bool upload)
{
    Slot& s = c->slots[sid];
    if (!c->inited) return fail(GSM_ESTATE, "init_data has not run");
    if (s.P < 1) return fail(GSM_ESTATE, "no parameters set on stream %d", sid);
    if (N < 0 || N > c->B) return fail(GSM_EINVAL, "N %d outside [0, %lld]", N, (long long)c->B);
    CK(cudaSetDevice(c->devid));
    int r = ensure_result_cap(c, s, s.P, N);
    if (r) return r;
    const int nbp = c->nb + 2;
    EvalArgs a{};
    a.bursts = c->d_bursts; a.pk = c->d_pk; a.pm = c->d_pm; a.nx = c->d_nx; a.cw = c->d_cw; a.ad = c->d_ad; a.dd = c->d_dd;
    a.params = s.d_params; a.ohist = c->d_ohist;
    a.whist = s.d_whist;
    // few vectors: the last block writes the result block straight into the pinned host mirror
    // (one launch, no copy on the stream); batches go through device memory and one DMA
    const bool direct = s.P <= kDirectResultMaxP;
    unsigned char* res = direct ? s.h_res_dev : s.d_res;
    a.hist_out = res_hist(res);
    a.chi2_out = res_chi2(res, s.P, nbp);
    a.flags_out = (unsigned long long*)res;
    a.counters = s.d_counters;
    a.mcE = (c->flags & GSM_KEEP_MCE) ? s.d_mcE : nullptr;
    a.sum_ai = (c->flags & GSM_KEEP_COUNTS) ? s.d_cnt : nullptr;
    a.n_steps = (c->flags & GSM_KEEP_COUNTS) ? s.d_steps : nullptr;
    a.b0 = 0; a.N = N; a.g_offset = c->g_offset;
    a.R = c->R; a.P = s.P; a.nb = c->nb; a.left_pad = c->left_pad;
    a.finalize = c->comm ? 0 : (c->peer_on ? 2 : 1);
    if (c->peer_on) {
        if (s.P > c->peer_cap_P) return fail(GSM_EINVAL, "%d vectors exceed the peer exchange capacity %d", s.P, c->peer_cap_P);
        if (nbp != c->peer_nbp) return fail(GSM_ESTATE, "histogram size changed after gsm_peer_export");
        for (int rk = 0; rk < c->peer_world; ++rk) a.peer.block[rk] = c->peer_map[rk] + (size_t)sid * c->peer_slot_words;
        a.peer.stride = (unsigned long long)c->peer_cap_P * nbp;
        a.peer.seq = s.peer_seq + 1;   // committed below, once the launch is in the stream
        a.peer.world = c->peer_world;
        a.peer.rank = c->rank;
        a.peer.timeout_ms = c->peer_timeout_ms;
    }
    a.seed = c->seed; a.epoch0 = s.epoch;
    a.epoch_per_vector = s.per_vector_epochs ? 1 : 0;
    a.clk_p = c->clk_p; a.bg_ad_rate = c->bg_ad; a.bg_dd_rate = c->bg_dd;

    if (c->mode == GSM_MODE_REFERENCE) {
        if (s.P != 1) return fail(GSM_EINVAL, "reference-draws mode evaluates one vector per launch (the reference's thread = one (burst, repeat))");
        if (c->peer_on || c->comm) return fail(GSM_ESTATE, "reference-draws mode does not combine with burst-split runs");
        const long long NN = (long long)N * c->R;
        if (NN > s.sob_cap) {
            CK(cudaStreamSynchronize(s.stream));
            if (s.d_sob_state) CK(cudaFree(s.d_sob_state));
            if (s.d_xs_state) CK(cudaFree(s.d_xs_state));
            s.d_sob_state = nullptr; s.d_xs_state = nullptr; s.sob_cap = 0;
            CK(cudaMalloc((void**)&s.d_sob_state, (size_t)NN * sizeof(ulonglong2)));
            CK(cudaMalloc((void**)&s.d_xs_state, (size_t)NN * sizeof(uint4)));
            s.sob_cap = NN;
        }
        if (NN != s.sob_NN) {   // gpuWorker calls init_randstate when N changes (gpuWorker.cpp:74-77)
            CK(launch_setup_ref(c->d_sob_dv, c->d_sob_sc, NN, s.d_sob_state, s.d_xs_state, s.stream));
            c->launches += NN > 0 ? 1 : 0;
            s.sob_NN = NN;
        }
        a.sob_dv = c->d_sob_dv; a.sob_state = s.d_sob_state; a.xs_state = s.d_xs_state;
    }
    const bool timing = !(c->flags & GSM_NO_TIMING);
    if (timing) CK(cudaEventRecord(s.ev0, s.stream));
    if (s.P == 1 && !c->comm) {
        // one vector: it rides in the launch parameters (run_resident re-sends the same bytes)
        a.use_inl = 1;
        a.inl = s.h_params[0];
        s.params_resident = true;
    } else if (upload) {
        CK(cudaMemcpyAsync(s.d_params, s.h_params, (size_t)s.P * sizeof(ParamRaw), cudaMemcpyHostToDevice, s.stream));
        s.params_resident = true;
    }
    // the dense build (48 warps per SM) where the walk is dominated by divergent loops and dependent
    // loads: background draws, or three and more states (pda_eval.cuh, PDA_MIN_BLOCKS_DENSE)
    // (measured, profiles/r02s_*, r02t_*: replay +9 % with background, +2 % with three states, -2 % on the
    // two-state walk; normal mode +12 % with background, -5 % / -11 % without)
    a.force_chunk_log2 = c->chunk_force; a.force_span = c->span_force; a.force_tail = c->tail_force;
    const bool dense = c->occ_force >= 0 ? c->occ_force != 0
                                         : (c->with_bg || (c->mode != GSM_MODE_NORMAL && s.h_params[0].n >= 3));
    CK(launch_eval(a, c->layout, c->with_bg, c->cfg, s.stream, c->mode, dense));
    c->launches += 1;
    if (c->peer_on) s.peer_seq = a.peer.seq;
    if (c->comm) {
        int e = g_nccl.AllReduce(s.d_whist, s.d_whist, (size_t)s.P * nbp, kNcclUint64, kNcclSum, c->comm, s.stream);
        if (e) return fail(GSM_ENCCL, "ncclAllReduce returned %s(%d)", g_nccl.GetErrorString ? g_nccl.GetErrorString(e) : "?", e);
        CK(launch_finalize(a, s.stream));
        c->launches += 1;
    }
    if (!direct) CK(cudaMemcpyAsync(s.h_res, s.d_res, res_bytes(s.P, nbp), cudaMemcpyDeviceToHost, s.stream));
    const size_t samples = (size_t)s.P * (size_t)N * (size_t)c->R;
    if (samples) {
        if (a.mcE) CK(cudaMemcpyAsync(s.h_mcE, s.d_mcE, samples * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        if (a.sum_ai) CK(cudaMemcpyAsync(s.h_cnt, s.d_cnt, samples * sizeof(long long), cudaMemcpyDeviceToHost, s.stream));
        if (a.n_steps) CK(cudaMemcpyAsync(s.h_steps, s.d_steps, samples * sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
    }
    if (timing) CK(cudaEventRecord(s.ev1, s.stream));
    s.res_P = s.P; s.res_nbp = nbp;
    if (!s.per_vector_epochs) s.epoch += (uint32_t)s.P;   // a coalesced launch advanced its members' epochs instead
    s.per_vector_epochs = false;
    s.alias_sid = -1;
    s.ran = true;
    return GSM_OK;
}

int gsm_run(gsm_ctx* c, int N, int sid)
{
    int r = check_sid(c, sid);
    if (r) return r;
    return run_common(c, N, sid, true);
}

int gsm_run_batch(gsm_ctx* c, int N, int sid, int n, int P, const float* args)
{
    int r = check_sid(c, sid);
    if (r) return r;
    if (P < 1 || !args) return fail(GSM_EINVAL, "batch of %d vectors", P);
    CK(cudaSetDevice(c->devid));
    Slot& s = c->slots[sid];
    r = ensure_params_cap(c, s, P);
    if (r) return r;
    CK(cudaStreamSynchronize(s.stream));
    const int na = n * n + n;
    for (int p = 0; p < P; ++p)
        if (!expand_params(n, args + (size_t)p * na, na, &s.h_params[p]))
            return fail(GSM_EINVAL, "vector %d: bad parameter vector for %d states", p, n);
    s.n = n;
    s.P = P;
    return run_common(c, N, sid, true);
}

// Several stream slots' single vectors in ONE launch on a carrier slot.  Vector p is evaluated
// with member p's own draw epoch (so the results are the bits the member's own launch would have
// produced), the member's epoch advances by one as if it had run, and until the member runs by
// itself again its result accessors (gsm_query, gsm_sync, gsm_result_chi2/hist/flags) read
// vector p of the carrier.  The carrier must not be a member, and must not be launched again
// before the members' results have been read.
int gsm_run_coalesced(gsm_ctx* c, int N, int carrier_sid, int n, int P, const float* args, const int* member_sids)
{
    int r = check_sid(c, carrier_sid);
    if (r) return r;
    if (P < 1 || !args || !member_sids) return fail(GSM_EINVAL, "coalesced launch of %d vectors", P);
    if (c->flags & (GSM_KEEP_MCE | GSM_KEEP_COUNTS)) return fail(GSM_ESTATE, "per-sample outputs are kept per stream: no coalesced launches");
    if (c->mode == GSM_MODE_REFERENCE) return fail(GSM_ESTATE, "reference-draws mode evaluates one vector per launch");
    for (int p = 0; p < P; ++p) {
        const int m = member_sids[p];
        if (m < 0 || m >= (int)c->slots.size() || m == carrier_sid) return fail(GSM_EINVAL, "member %d: stream %d (carrier %d)", p, m, carrier_sid);
        for (int k = 0; k < p; ++k)
            if (member_sids[k] == m) return fail(GSM_EINVAL, "stream %d is a member twice", m);
    }
    CK(cudaSetDevice(c->devid));
    Slot& s = c->slots[carrier_sid];
    r = ensure_params_cap(c, s, P);
    if (r) return r;
    CK(cudaStreamSynchronize(s.stream));
    const int na = n * n + n;
    for (int p = 0; p < P; ++p) {
        if (!expand_params(n, args + (size_t)p * na, na, &s.h_params[p]))
            return fail(GSM_EINVAL, "vector %d: bad parameter vector for %d states", p, n);
        s.h_params[p].epoch = c->slots[member_sids[p]].epoch;
    }
    s.n = n;
    s.P = P;
    s.per_vector_epochs = true;
    r = run_common(c, N, carrier_sid, true);
    if (r) { s.per_vector_epochs = false; return r; }
    for (int p = 0; p < P; ++p) {
        Slot& m = c->slots[member_sids[p]];
        m.epoch += 1u;
        m.alias_sid = carrier_sid;
        m.alias_p = p;
        m.n = n;
    }
    return GSM_OK;
}

int gsm_run_resident(gsm_ctx* c, int N, int sid)
{
    int r = check_sid(c, sid);
    if (r) return r;
    if (!c->slots[sid].params_resident) return fail(GSM_ESTATE, "no resident parameters on stream %d", sid);
    return run_common(c, N, sid, false);
}

// where the results of stream sid live: itself, or the carrier of the coalesced launch it was part of
static inline const Slot& result_slot(const gsm_ctx* c, int sid, int* p_base)
{
    const Slot& s = c->slots[sid];
    if (s.alias_sid >= 0) { *p_base = s.alias_p; return c->slots[s.alias_sid]; }
    *p_base = 0;
    return s;
}

int gsm_query(gsm_ctx* c, int sid)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size()) return 0;  // mc::streamQuery returns false
    int pb;
    cudaError_t e = cudaStreamQuery(result_slot(c, sid, &pb).stream);
    if (e == cudaSuccess) return 1;
    if (e == cudaErrorNotReady) { cudaGetLastError(); return 0; }
    return fail(GSM_ECUDA, "cudaStreamQuery returned %s(%d)", cudaGetErrorString(e), (int)e);
}

int gsm_sync(gsm_ctx* c, int sid)
{
    int r = check_sid(c, sid);
    if (r) return r;
    int pb;
    CK(cudaStreamSynchronize(result_slot(c, sid, &pb).stream));
    return GSM_OK;
}

float gsm_result_chi2(const gsm_ctx* c, int sid)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size()) return NAN;
    int pb;
    const Slot& s = result_slot(c, sid, &pb);
    if (!s.ran || pb >= s.res_P) return NAN;
    return res_chi2(s.h_res, s.res_P, s.res_nbp)[pb];
}
const float* gsm_result_chi2v(const gsm_ctx* c, int sid, int* P)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size()) return nullptr;
    int pb;
    const Slot& s = result_slot(c, sid, &pb);
    if (!s.ran || pb >= s.res_P) return nullptr;
    if (P) *P = c->slots[sid].alias_sid >= 0 ? 1 : s.res_P;
    return res_chi2(s.h_res, s.res_P, s.res_nbp) + pb;
}
const uint64_t* gsm_result_hist(const gsm_ctx* c, int sid, int p)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size()) return nullptr;
    int pb;
    const Slot& s = result_slot(c, sid, &pb);
    if (!s.ran) return nullptr;
    if (c->slots[sid].alias_sid >= 0 && p != 0) return nullptr;   // a member owns one vector of the carrier
    p += pb;
    if (p < 0 || p >= s.res_P) return nullptr;
    return (const uint64_t*)(res_hist(s.h_res) + (size_t)p * s.res_nbp);
}
const float* gsm_result_mcE(const gsm_ctx* c, int sid)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size()) return nullptr;
    return c->slots[sid].h_mcE;
}
const int64_t* gsm_result_counts(const gsm_ctx* c, int sid)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size()) return nullptr;
    return (const int64_t*)c->slots[sid].h_cnt;
}
const int32_t* gsm_result_steps(const gsm_ctx* c, int sid)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size()) return nullptr;
    return c->slots[sid].h_steps;
}
int gsm_result_timing(const gsm_ctx* c, int sid, uint64_t out_ns[4])
{
    if (!c || !out_ns || sid < 0 || sid >= (int)c->slots.size()) return fail(GSM_EINVAL, "bad timing query");
    int pb;
    const Slot& s = result_slot(c, sid, &pb);   // a member of a coalesced launch reads its carrier's stamps
    if (!s.ran) return fail(GSM_ESTATE, "no run on stream %d yet", sid);
    const ResultHeader* h = (const ResultHeader*)s.h_res;
    for (int i = 0; i < 4; ++i) out_ns[i] = h->stamp[i];
    return GSM_OK;
}

uint64_t gsm_result_flags(const gsm_ctx* c, int sid)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size()) return 0;
    int pb;
    const Slot& s = result_slot(c, sid, &pb);
    return s.ran ? ((const ResultHeader*)s.h_res)->flags : 0;
}
float gsm_last_elapsed_ms(gsm_ctx* c, int sid)
{
    if (!c || sid < 0 || sid >= (int)c->slots.size() || !c->slots[sid].ran || (c->flags & GSM_NO_TIMING)) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, c->slots[sid].ev0, c->slots[sid].ev1) != cudaSuccess) { cudaGetLastError(); return -1.0f; }
    return ms;
}
uint64_t gsm_launch_count(const gsm_ctx* c) { return c ? c->launches.load() : 0; }

int gsm_comm_unique_id(void* out128)
{
    if (!out128) return fail(GSM_EINVAL, "null id");
    int r = nccl_load();
    if (r) return r;
    int e = g_nccl.GetUniqueId(out128);
    if (e) return fail(GSM_ENCCL, "ncclGetUniqueId returned %d", e);
    return GSM_OK;
}

int gsm_comm_init(gsm_ctx* c, const void* id128, int rank, int world)
{
    if (!c || !id128 || world < 1 || rank < 0 || rank >= world) return fail(GSM_EINVAL, "rank %d of %d", rank, world);
    if (c->peer_on) return fail(GSM_ESTATE, "the peer exchange is connected: use one reduction or the other");
    int r = nccl_load();
    if (r) return r;
    CK(cudaSetDevice(c->devid));
    Id128 id;
    memcpy(id.b, id128, 128);
    int e = g_nccl.CommInitRank(&c->comm, world, id, rank);
    if (e) {
        c->comm = nullptr;
        return fail(GSM_ENCCL, "ncclCommInitRank returned %s(%d)", g_nccl.GetErrorString ? g_nccl.GetErrorString(e) : "?", e);
    }
    c->rank = rank;
    c->world = world;
    return GSM_OK;
}

int gsm_comm_destroy(gsm_ctx* c)
{
    if (!c) return fail(GSM_EINVAL, "null context");
    if (c->comm) {
        CK(cudaSetDevice(c->devid));
        for (Slot& s : c->slots) CK(cudaStreamSynchronize(s.stream));
        g_nccl.CommDestroy(c->comm);
        c->comm = nullptr;
    }
    c->rank = 0; c->world = 1;
    return GSM_OK;
}

int gsm_peer_export(gsm_ctx* c, int world, int max_P, void* handle_out)
{
    if (!c || !handle_out || world < 1 || world > kMaxPeers || max_P < 1)
        return fail(GSM_EINVAL, "peer export: world %d (1..%d), max_P %d", world, kMaxPeers, max_P);
    if (c->comm) return fail(GSM_ESTATE, "an NCCL communicator is attached: use one reduction or the other");
    // peers may have this block mapped and may be storing into it: it is only replaced after a disconnect
    if (c->peer_on) return fail(GSM_ESTATE, "the peer exchange is connected: gsm_peer_disconnect first");
    CK(cudaSetDevice(c->devid));
    for (Slot& s : c->slots) CK(cudaStreamSynchronize(s.stream));
    peer_release(c);
    const int nbp = c->nb + 2;
    c->peer_slot_words = kPeerFlagWords + (size_t)2 * world * max_P * nbp;
    c->peer_bytes = c->slots.size() * c->peer_slot_words * sizeof(unsigned long long);
    CK(cudaMalloc((void**)&c->peer_base, c->peer_bytes));
    CK(cudaMemset(c->peer_base, 0, c->peer_bytes));
    c->peer_cap_P = max_P; c->peer_nbp = nbp; c->peer_world = world;
    for (Slot& s : c->slots) s.peer_seq = 0;
    PeerHandle h;
    memset(&h, 0, sizeof(h));
    h.magic = kPeerMagic; h.pid = (uint32_t)getpid(); h.devid = c->devid; h.world = world;
    h.ptr = (uint64_t)(uintptr_t)c->peer_base; h.bytes = c->peer_bytes;
    for (int i = 0; i < 16; ++i) {
        auto nib = [](char ch) { return ch <= '9' ? ch - '0' : (ch | 0x20) - 'a' + 10; };
        h.uuid[i] = (unsigned char)((nib(c->uuid[2 * i]) << 4) | nib(c->uuid[2 * i + 1]));
    }
    CK(cudaIpcGetMemHandle(&h.ipc, c->peer_base));
    memcpy(handle_out, &h, sizeof(h));
    return GSM_OK;
}

int gsm_peer_connect(gsm_ctx* c, const void* handles, int rank, int world)
{
    if (!c || !handles || world < 1 || world > kMaxPeers || rank < 0 || rank >= world)
        return fail(GSM_EINVAL, "peer connect: rank %d of %d", rank, world);
    if (!c->peer_base || c->peer_world != world) return fail(GSM_ESTATE, "gsm_peer_export(world = %d) has not run", world);
    CK(cudaSetDevice(c->devid));
    const PeerHandle* hs = (const PeerHandle*)handles;
    {
        // Forward progress (include/gsmfret_c.h): ranks that share this GPU wait for each other inside
        // their kernels, which only works while the GPU runs them concurrently.
        PeerHandle me;
        memcpy(&me, hs + rank, sizeof(me));
        int same_gpu = 0, same_gpu_other_process = 0;
        for (int r = 0; r < world; ++r) {
            PeerHandle h;
            memcpy(&h, hs + r, sizeof(h));
            if (memcmp(h.uuid, me.uuid, 16) == 0) {
                same_gpu++;
                if (h.pid != me.pid) same_gpu_other_process++;
            }
        }
        if (same_gpu > 1) {
            if (!(c->flags & GSM_PEER_SHARED_DEVICE))
                return fail(GSM_ESTATE, "%d ranks of this exchange share GPU %d: their kernels wait for each other and CUDA does "
                            "not promise to run them concurrently; use one rank per GPU, gsm_comm_init, or create the "
                            "contexts with GSM_PEER_SHARED_DEVICE", same_gpu, c->devid);
            // ranks of other processes on this GPU: progress then rests on the driver's time-slicing
            // between contexts (or MPS) — slow but live; the caller vouched for it with the flag
            (void)same_gpu_other_process;
            const char* lb = getenv("CUDA_LAUNCH_BLOCKING");
            if (lb && atoi(lb) != 0) return fail(GSM_ESTATE, "CUDA_LAUNCH_BLOCKING serialises the ranks that share GPU %d", c->devid);
            const char* mc = getenv("CUDA_DEVICE_MAX_CONNECTIONS");
            const int queues = mc && atoi(mc) > 0 ? atoi(mc) : 8;
            if ((long long)same_gpu * (long long)c->slots.size() > queues)
                return fail(GSM_ESTATE, "%d ranks x %d stream slots on GPU %d exceed its %d hardware queues "
                            "(CUDA_DEVICE_MAX_CONNECTIONS): launches would queue behind the kernels that wait for them",
                            same_gpu, (int)c->slots.size(), c->devid, queues);
        }
    }
    for (int r = 0; r < world; ++r) {
        PeerHandle h;
        memcpy(&h, hs + r, sizeof(h));
        if (h.magic != kPeerMagic || h.world != world || h.bytes != c->peer_bytes)
            return fail(GSM_EINVAL, "peer handle %d does not match (world %d, %llu bytes)", r, h.world, (unsigned long long)h.bytes);
        if (r == rank) {
            if (h.ptr != (uint64_t)(uintptr_t)c->peer_base || h.pid != (uint32_t)getpid())
                return fail(GSM_EINVAL, "handle %d is not this context's export", r);
            c->peer_map[r] = c->peer_base;
        } else if (h.pid == (uint32_t)getpid()) {
            // another context of this process: the pointer is already valid here
            if (h.devid != c->devid) {
                cudaError_t e = cudaDeviceEnablePeerAccess(h.devid, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                    return fail(GSM_ECUDA, "cudaDeviceEnablePeerAccess(%d) returned %s(%d)", h.devid, cudaGetErrorString(e), (int)e);
                cudaGetLastError();
            }
            c->peer_map[r] = (unsigned long long*)(uintptr_t)h.ptr;
        } else {
            void* m = nullptr;
            CK(cudaIpcOpenMemHandle(&m, h.ipc, cudaIpcMemLazyEnablePeerAccess));
            c->peer_map[r] = (unsigned long long*)m;
            c->peer_ipc[r] = true;
        }
    }
    c->rank = rank; c->world = world;
    c->peer_on = world > 1;
    return GSM_OK;
}

int gsm_peer_disconnect(gsm_ctx* c)
{
    if (!c) return fail(GSM_EINVAL, "null context");
    CK(cudaSetDevice(c->devid));
    for (Slot& s : c->slots) CK(cudaStreamSynchronize(s.stream));
    peer_release(c);
    c->rank = 0; c->world = 1;
    return GSM_OK;
}

int gsm_peer_set_timeout(gsm_ctx* c, uint32_t ms)
{
    if (!c || ms == 0) return fail(GSM_EINVAL, "timeout of %u ms", ms);
    c->peer_timeout_ms = ms;
    return GSM_OK;
}

int gsm_launch_shape(int64_t N, int R, int P, int resident_blocks, int guided, int64_t* out6)
{
    if (N < 0 || R < 1 || P < 1 || resident_blocks < 1 || !out6) return fail(GSM_EINVAL, "bad launch shape query");
    EvalArgs a{};
    a.N = N; a.R = R; a.P = P; a.force_tail = -1;
    const int blocks = eval_launch_shape(a, resident_blocks, guided != 0);
    out6[0] = a.chunk; out6[1] = a.span; out6[2] = a.tail_log2; out6[3] = a.big_chunks; out6[4] = a.chunks_per_p;
    out6[5] = blocks;
    return GSM_OK;
}

int gsm_issue_peak(gsm_ctx* c, int kind, double* out)
{
    if (!c || !out || (kind != 0 && kind != 1)) return fail(GSM_EINVAL, "bad issue-peak call");
    CK(cudaSetDevice(c->devid));
    int sms = 148;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->devid));
    unsigned int* d = nullptr;
    CK(cudaMalloc(&d, sizeof(unsigned int)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int grid = sms * 8, iters = 1 << 15;   // 8 x 256 threads per SM: every scheduler busy
    double best = 0.0;
    cudaError_t e = cudaSuccess;
    for (int rep = 0; rep < 4 && e == cudaSuccess; ++rep) {   // first pass warms the clocks up
        cudaEventRecord(e0, 0);
        e = launch_issue_probe(d, grid, iters, kind == 0, 0);
        cudaEventRecord(e1, 0);
        if (e == cudaSuccess) e = cudaEventSynchronize(e1);
        float ms = 0.0f;
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
        if (e == cudaSuccess && ms > 0.0f) {
            const double ops = (double)grid * 256.0 * (double)iters * 16.0 / ((double)ms * 1e-3);
            if (ops > best) best = ops;
        }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    CK(e);
    *out = best;
    return GSM_OK;
}

int gsm_probe(gsm_ctx* c, int what, int64_t n, int ni, int nj, const uint32_t* w0,
              const uint32_t* w1, const float* f0, const float* f1, const double* d0,
              const int32_t* i0, uint32_t* ou, float* of, double* od, int64_t* ol, int32_t* oi)
{
    if (!c || n < 0) return fail(GSM_EINVAL, "bad probe call");
    CK(cudaSetDevice(c->devid));
    // element counts per input/output by probe kind
    size_t nw0 = 0, nw1 = 0, nf0 = 0, nf1 = 0, nd0 = 0, ni0 = 0, nou = 0, nof = 0, nod = 0, nol = 0, noi = 0;
    const size_t N = (size_t)n;
    switch (what) {
    case 0: nw0 = 4 * N; nw1 = 2 * N; nou = 4 * N; break;
    case 1: case 5: nf0 = N; nof = N; break;
    case 2: case 3: nd0 = N; nod = N; break;
    case 4: case 6: nw0 = N; nof = N; break;
    case 7: nw0 = 2 * N; ni0 = N; nd0 = N; nol = N; break;
    case 8: nw0 = 2 * N; nd0 = N; nou = N; break;
    case 9: nw0 = N; nf0 = N; nf1 = N; nof = N; break;
    case 10: nw0 = N; nf0 = N; nof = N; break;
    case 11: if (ni < 1 || ni > 16) return fail(GSM_EINVAL, "classes %d", ni); nw0 = 2 * N; nf0 = ni; noi = N * ni; break;
    case 12: if (ni < 1 || ni > 16) return fail(GSM_EINVAL, "classes %d", ni); nw0 = N; nf0 = ni; noi = N; break;
    case 13: nf0 = N; noi = N; break;
    case 14: nd0 = 2 * N; nod = N; break;
    default: return fail(GSM_EINVAL, "probe kind %d", what);
    }
    ProbeArgs a{};
    a.what = what; a.n = n; a.ni = ni; a.nj = nj;
    std::vector<void*> tofree;
    auto up = [&](const void* h, size_t bytes, const void** d) -> cudaError_t {
        *d = nullptr;
        if (!bytes) return cudaSuccess;
        if (!h) return cudaErrorInvalidValue;
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) return e;
        tofree.push_back(p);
        *d = p;
        return cudaMemcpy(p, h, bytes, cudaMemcpyHostToDevice);
    };
    auto mk = [&](void* h, size_t bytes, void** d) -> cudaError_t {
        *d = nullptr;
        if (!bytes) return cudaSuccess;
        if (!h) return cudaErrorInvalidValue;
        cudaError_t e = cudaMalloc(d, bytes);
        if (e == cudaSuccess) tofree.push_back(*d);
        return e;
    };
    cudaError_t e = cudaSuccess;
    if (e == cudaSuccess) e = up(w0, nw0 * 4, (const void**)&a.w0);
    if (e == cudaSuccess) e = up(w1, nw1 * 4, (const void**)&a.w1);
    if (e == cudaSuccess) e = up(f0, nf0 * 4, (const void**)&a.f0);
    if (e == cudaSuccess) e = up(f1, nf1 * 4, (const void**)&a.f1);
    if (e == cudaSuccess) e = up(d0, nd0 * 8, (const void**)&a.d0);
    if (e == cudaSuccess) e = up(i0, ni0 * 4, (const void**)&a.i0);
    if (e == cudaSuccess) e = mk(ou, nou * 4, (void**)&a.ou);
    if (e == cudaSuccess) e = mk(of, nof * 4, (void**)&a.of);
    if (e == cudaSuccess) e = mk(od, nod * 8, (void**)&a.od);
    if (e == cudaSuccess) e = mk(ol, nol * 8, (void**)&a.ol);
    if (e == cudaSuccess) e = mk(oi, noi * 4, (void**)&a.oi);
    if (e == cudaSuccess) e = launch_probe(a, 0);
    if (e == cudaSuccess) { c->launches += (n ? 1 : 0); e = cudaDeviceSynchronize(); }
    if (e == cudaSuccess && nou) e = cudaMemcpy(ou, a.ou, nou * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && nof) e = cudaMemcpy(of, a.of, nof * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && nod) e = cudaMemcpy(od, a.od, nod * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && nol) e = cudaMemcpy(ol, a.ol, nol * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && noi) e = cudaMemcpy(oi, a.oi, noi * 4, cudaMemcpyDeviceToHost);
    for (void* p : tofree) cudaFree(p);
    if (e != cudaSuccess) return fail(GSM_ECUDA, "probe %d returned %s(%d)", what, cudaGetErrorString(e), (int)e);
    return GSM_OK;
}

}  // extern "C"
