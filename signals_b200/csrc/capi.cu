// capi.cu -- host runtime behind the C ABI of include/signals_b200.h.
//
// One context per CUDA device (created on first use): two slot streams for the chunked
// host->device->host pipeline, grow-only device buffers, the per-call constant tables.
// The host never touches the per-bin data: validation, narrowing and re-layout happen in k_pack.
// There is no CPU execution path here -- if CUDA is unusable every entry point fails loudly.
#include "../../include/signals_b200.h"
#include "kernels.h"
#include "host_stage.h"
#include <algorithm>
#include <memory>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

thread_local std::string t_err;

namespace {

int fail(const std::string& msg)
{
    t_err = msg;
    return 1;
}

#define SGB_CUDA(call)                                                                                       \
    do {                                                                                                     \
        cudaError_t e__ = (call);                                                                            \
        if (e__ != cudaSuccess) {                                                                            \
            char b__[512];                                                                                   \
            snprintf(b__, sizeof b__, "CUDA error %s at %s:%d: %s", cudaGetErrorName(e__), __FILE__, __LINE__, \
                     cudaGetErrorString(e__));                                                               \
            return fail(b__);                                                                                \
        }                                                                                                    \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) return 0;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            p = nullptr;
            return fail(std::string("cudaMalloc of ") + std::to_string(want) + " bytes failed: " + cudaGetErrorString(e));
        }
        cap = want;
        return 0;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
    }
};

// grow-only pinned host buffer: plan arrays are uploaded from page-locked memory so that the copies are truly
// asynchronous (a cudaMemcpyAsync from pageable memory first synchronises the stream with the host, which
// would stall the chunk pipeline once per chunk)
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) return 0;
        if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 4 + 4096;
        cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocDefault);
        if (e != cudaSuccess) {
            p = nullptr;
            return fail(std::string("cudaHostAlloc of ") + std::to_string(want) + " bytes failed: " + cudaGetErrorString(e));
        }
        cap = want;
        return 0;
    }
    void release()
    {
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
    }
};

// device scratch of the split schedule (k_emit + k_recur) for one plan / one chunk slot: grow-only
struct SplitWork {
    DevBuf cls, perm, part, g_npl, g_tmax, g_eoff, g_size, misc, E;
    int reserve(const SgbSplitSizes& z)
    {
        return cls.reserve(z.cls) || perm.reserve(z.perm) || part.reserve(z.part) || g_npl.reserve(z.g_i32) ||
               g_tmax.reserve(z.g_i32) || g_eoff.reserve(z.g_i64) || g_size.reserve(z.g_i64) || misc.reserve(64) ||
               E.reserve(z.e_bytes);
    }
    SgbSplitDev dev(const SgbSplitSizes& z) const
    {
        SgbSplitDev W;
        W.cls = (uint8_t*)cls.p; W.perm = (int*)perm.p; W.part = (unsigned int*)part.p; W.g_npl = (int*)g_npl.p;
        W.g_tmax = (int*)g_tmax.p; W.g_eoff = (long long*)g_eoff.p; W.g_size = (unsigned long long*)g_size.p;
        W.ngroups = (int*)misc.p; W.counters = (unsigned int*)misc.p + 4;
        W.E = (double*)E.p; W.e_cap_sectors = (long long)(z.e_bytes / 32); W.gcap = z.gcap; W.ntile = z.ntile;
        return W;
    }
    size_t bytes() const
    {
        return cls.cap + perm.cap + part.cap + g_npl.cap + g_tmax.cap + g_eoff.cap + g_size.cap + misc.cap + E.cap;
    }
    void release()
    {
        cls.release(); perm.release(); part.release(); g_npl.release(); g_tmax.release(); g_eoff.release();
        g_size.release(); misc.release(); E.release();
    }
};

// Slices of a length-descending plan for the split schedule: each [q0, q1) range of plan entries keeps the emission
// tiles under `budget` bytes (the whole batch is one slice unless it is very large).
struct SplitSlice {
    long long q0, q1, rows;
    int maxlen;
};
std::vector<SplitSlice> split_slices(const std::vector<int>& len, int S, size_t budget)
{
    std::vector<SplitSlice> out;
    const long long n = (long long)len.size();
    long long q0 = 0;
    while (q0 < n) {
        long long q1 = q0, rows = 0;
        const int maxlen = len[(size_t)q0];              // entries are sorted longest first (padding entries have length 0)
        int mx = maxlen;
        while (q1 < n) {
            const long long step = std::min<long long>(n - q1, 32 * 64);
            long long add = 0;
            for (long long i = q1; i < q1 + step; i++) { add += len[(size_t)i]; if (len[(size_t)i] > mx) mx = len[(size_t)i]; }
            if (q1 > q0 && sgbk_split_sizes(q1 + step - q0, rows + add, mx, S).e_bytes > budget) break;
            rows += add;
            q1 += step;
        }
        out.push_back({q0, q1, rows, mx});
        q0 = q1;
    }
    return out;
}

size_t split_budget_bytes()
{
    const char* e = getenv("SGB_E_BUDGET_MB");
    long long mb = e ? atoll(e) : 0;
    if (mb <= 0) mb = 8192;
    return (size_t)mb << 20;
}

// Which schedule a decode = reference call takes: 1 = split (k_emit + k_recur), 0 = fused k_decode.
// The fused kernel has one lane per sequence and needs ~75,000 sequences to fill a B200; below that the split schedule
// (emission parallel over bins) is up to 1.7x faster, above it the two are within a few per cent and the fused kernel
// needs no emission scratch.  SGB_SCHED=split|fused forces one; SGB_SPLIT_MAX_SEQ moves the switch-over point.
int want_split(const SgbModelArgs& m, long long nseq)
{
    if (m.decode != SGB_DECODE_REFERENCE) return 0;          // chained decode keeps the fused kernel + k_backtrace
    const char* e = getenv("SGB_SCHED");
    if (e && !strcmp(e, "fused")) return 0;
    if (e && !strcmp(e, "split")) return 1;
    long long lim = 60000;
    const char* t = getenv("SGB_SPLIT_MAX_SEQ");
    if (t && atoll(t) > 0) lim = atoll(t);
    // beta-binomial: the emission dominates and the split kernels evaluate it faster at every batch size
    // (10,000 cells ASCN: 3.78 ms against the fused kernel's 3.86 ms); binomial: the fused kernel wins once the device is full
    return m.lik_bb || nseq < lim;
}

// host-side description of a batch of sequences, sorted by length (longest first)
struct HostPlan {
    std::vector<long long> goff, ostart;       // goff: per group of 32 sequences, ostart: per sequence
    std::vector<int> len;
    long long packed_len = 0;   // records
    long long nseq = 0;
    long long rows = 0;         // total bins covered
};

// seg_off[s0..s1]; row offsets are made relative to `row_base`
int build_host_plan(const int64_t* seg_off, int64_t s0, int64_t s1, int64_t row_base, HostPlan& hp, long long warp_slots,
                    int group_hint)
{
    const int64_t nseg = s1 - s0;
    hp.goff.clear(); hp.ostart.clear(); hp.len.clear();
    hp.packed_len = 0; hp.nseq = 0; hp.rows = 0;
    int64_t maxlen = 0;
    for (int64_t s = s0; s < s1; s++) {
        int64_t L = seg_off[s + 1] - seg_off[s];
        if (L < 0) return fail("seg_off must be non-decreasing");
        if (L > 0x7fffffff) return fail("a sequence longer than 2^31-1 bins is not supported");
        if (L > maxlen) maxlen = L;
    }
    std::vector<int64_t> order;
    order.reserve((size_t)nseg);
    if (maxlen <= (1 << 22)) {
        // counting sort by length, descending, stable
        std::vector<int64_t> cnt((size_t)maxlen + 2, 0);
        for (int64_t s = s0; s < s1; s++) {
            int64_t L = seg_off[s + 1] - seg_off[s];
            if (L > 0) cnt[(size_t)(maxlen - L) + 1]++;
        }
        for (size_t i = 1; i < cnt.size(); i++) cnt[i] += cnt[i - 1];
        order.resize((size_t)cnt[cnt.size() - 1]);
        for (int64_t s = s0; s < s1; s++) {
            int64_t L = seg_off[s + 1] - seg_off[s];
            if (L > 0) order[(size_t)cnt[(size_t)(maxlen - L)]++] = s;
        }
    } else {
        for (int64_t s = s0; s < s1; s++) if (seg_off[s + 1] > seg_off[s]) order.push_back(s);
        std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) {
            return (seg_off[a + 1] - seg_off[a]) > (seg_off[b + 1] - seg_off[b]);
        });
    }
    // Sequences per warp.  32 fills every lane; when the batch is too small to give every resident warp of
    // the device a group, narrower groups (lanes beyond G carry empty sequences) trade lane utilisation for
    // more, shorter-latency warps.  The kernels need no change: an empty lane is just a sequence of length 0.
    const long long nreal = (long long)order.size();
    int G = 32;
    if (group_hint > 0) G = group_hint;
    else if (warp_slots > 0) {
        if ((nreal + 31) / 32 < warp_slots) G = 16;       // measured on 1,000 cells x 5,700 bins: 16 helps ~12 %,
        if ((nreal + 31) / 32 < warp_slots / 2) G = 8;    // 8 another ~5 %, 4 loses
    }
    const long long ngroups = (nreal + G - 1) / G;
    hp.nseq = ngroups * 32;                       // padded: kernels index sequence r = group * 32 + lane
    hp.ostart.assign((size_t)hp.nseq, 0);
    hp.len.assign((size_t)hp.nseq, 0);
    hp.goff.resize((size_t)ngroups);
    long long p = 0;
    for (long long i = 0; i < nreal; i++) {
        int64_t s = order[(size_t)i];
        int L = (int)(seg_off[s + 1] - seg_off[s]);
        const long long g = i / G, lane = i % G, r = g * 32 + lane;
        hp.ostart[(size_t)r] = (long long)(seg_off[s] - row_base);
        hp.len[(size_t)r] = L;
        hp.rows += L;
        if (lane == 0) {
            // the first sequence of a group is its longest; tile = roundup(T, SGB_CH) bins x 32 lanes
            hp.goff[(size_t)g] = p;
            p += (((long long)L + SGB_CH - 1) / SGB_CH) * SGB_CH * 32;
        }
    }
    hp.packed_len = p;
    return 0;
}

struct Slot {
    cudaStream_t st = nullptr;
    DevBuf n, x, cn, cn8, out, packed, bp, last_state, goff, ostart, len, loglik;
    SplitWork split;
    unsigned int* counter = nullptr;
    HostPlan hp;
    PinBuf stage_in, stage_out;   // page-locked staging of the chunk's (narrowed) columns / decoded states
    int64_t pend_r0 = 0, pend_rows = 0;      // decoded states waiting in stage_out for the copy into the caller's buffer
    PinBuf pin;                 // page-locked copy of hp's arrays for the upload
    // pending device->host copy bookkeeping is implicit: copies are enqueued on st
};

struct Ctx {
    int device = -1;
    int sm_count = 148;
    int group_hint = 0;         // SGB_GROUP=32|16|8 forces the sequences-per-warp choice (0 = automatic)
    int stage = 2;              // input staging of k_decode: 2 = RAW (default: caller's columns read in place, no k_pack),
                                // 1 = packed records by TMA bulk copies (SGB_STAGE=tma), 0 = packed records by LDG (SGB_STAGE=ldg)
    Slot slot[2];
    cudaStream_t main = nullptr;
    cudaEvent_t ev_setup = nullptr;
    SgbDevTables* d_tabs = nullptr;
    int* d_err = nullptr;
    DevBuf scratch_a, scratch_b, scratch_c, scratch_d;   // small utilities (viterbi, logspace_add, ...)
    std::unique_ptr<SgbPool> pool;                       // host staging workers (created on first use)
};

std::mutex g_mu;
Ctx* g_ctx[64] = {nullptr};

int get_ctx(Ctx** out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return fail(std::string("no usable CUDA device (this library has no CPU path): ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return fail("device index out of range");
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_ctx[dev]) {
        Ctx* c = new Ctx();
        c->device = dev;
        cudaDeviceProp prop;
        SGB_CUDA(cudaGetDeviceProperties(&prop, dev));
        if (prop.major < 10) {
            delete c;
            return fail(std::string("device ") + prop.name + " is not sm_100-class; libsignals_b200 is built for sm_100a only");
        }
        c->sm_count = prop.multiProcessorCount;
        {
            const char* e = getenv("SGB_STAGE");
            c->stage = !e ? 2 : (!strcmp(e, "ldg") || !strcmp(e, "0")) ? 0 : (!strcmp(e, "tma") || !strcmp(e, "1")) ? 1 : 2;
            const char* gsz = getenv("SGB_GROUP");
            if (gsz) { int v = atoi(gsz); if (v == 32 || v == 16 || v == 8 || v == 4) c->group_hint = v; }
        }
        SGB_CUDA(cudaStreamCreate(&c->main));
        SGB_CUDA(cudaEventCreateWithFlags(&c->ev_setup, cudaEventDisableTiming));
        for (int i = 0; i < 2; i++) {
            SGB_CUDA(cudaStreamCreate(&c->slot[i].st));
            SGB_CUDA(cudaMalloc(&c->slot[i].counter, sizeof(unsigned int)));
        }
        SGB_CUDA(cudaMalloc(&c->d_tabs, sizeof(SgbDevTables)));
        SGB_CUDA(cudaMalloc(&c->d_err, sizeof(int)));
        g_ctx[dev] = c;
    }
    *out = g_ctx[dev];
    return 0;
}

int check_params(const sgb_params* P, SgbModelArgs& m)
{
    if (!P) return fail("params is NULL");
    if (P->model != SGB_MODEL_HSCN && P->model != SGB_MODEL_ASCN) return fail("params.model must be SGB_MODEL_HSCN or SGB_MODEL_ASCN");
    if (P->likelihood != SGB_LIK_BINOMIAL && P->likelihood != SGB_LIK_BETABINOMIAL)
        return fail("Likelihood model must be one of binomial or betabinomial");      // R/callHSCN.R:778-780
    if (P->n_states < 1 || P->n_states > SGB_MAX_STATES) return fail("n_states (maxCN+1) must be in 1..32");
    if (P->decode != SGB_DECODE_REFERENCE && P->decode != SGB_DECODE_CHAINED) return fail("params.decode invalid");
    if (!(P->selftransitionprob >= 0.0 && P->selftransitionprob <= 1.0)) return fail("selftransitionprob must be in [0,1]");
    if (P->loherror != P->loherror || P->eps != P->eps || P->rho != P->rho || P->Abias != P->Abias)
        return fail("NaN parameter");
    m.model = P->model;
    m.lik_bb = (P->likelihood == SGB_LIK_BETABINOMIAL && P->rho > 0.0) ? 1 : 0;   // rho == 0: VGAM falls back to dbinom
    m.S = P->n_states;
    m.decode = P->decode;
    m.loherror = P->loherror;
    m.rho = P->rho;
    return 0;
}

const char* err_bits_text(int bits)
{
    if (bits & SGB_ERR_NEGATIVE_COUNT) return "negative read count (or NA_integer_) in totalcounts / alleleB";
    if (bits & SGB_ERR_COUNT_TOO_LARGE) return "read count >= 2^28 is not supported";
    if (bits & SGB_ERR_BAD_CN) return "copy-number state outside 0..255";
    if (bits & SGB_ERR_NOT_INTEGER) return "a numeric (double) column holds NA / NaN or a value that is not an integer";
    if (bits & SGB_ERR_INTERNAL) return "internal error: emission tile bound exceeded";
    if (bits & SGB_ERR_BAD_KEY) return "cell / chromosome code outside its declared range, negative start, or state outside 0..255";
    if (bits & SGB_ERR_KEY_CONFLICT) return "rows with the same (cell, chr, start) disagree on `state`";
    return "unknown input error";
}

// split schedule over a whole (length-descending) plan: slices, scratch reservation, launches
int run_split(SplitWork& W, std::vector<SplitSlice>& slices, const HostPlan& hp, const SgbPlanDev& pd0, const SgbColsIn& in,
              const SgbDevTables* d_tabs, const SgbModelArgs& m, uint8_t* d_state_out, int sm_count, cudaStream_t st,
              cudaEvent_t ev_prep, cudaEvent_t ev_mid)
{
    if (slices.empty()) slices = split_slices(hp.len, m.S, split_budget_bytes());
    SgbSplitSizes zmax = sgbk_split_sizes(0, 0, 0, m.S);
    for (const SplitSlice& sl : slices) {
        const SgbSplitSizes z = sgbk_split_sizes(sl.q1 - sl.q0, sl.rows, sl.maxlen, m.S);
        zmax.cls = std::max(zmax.cls, z.cls); zmax.perm = std::max(zmax.perm, z.perm); zmax.part = std::max(zmax.part, z.part);
        zmax.g_i32 = std::max(zmax.g_i32, z.g_i32); zmax.g_i64 = std::max(zmax.g_i64, z.g_i64);
        zmax.e_bytes = std::max(zmax.e_bytes, z.e_bytes);
    }
    if (W.reserve(zmax)) return 1;           // grow-only: allocates on the first call of a plan / slot
    bool first = true;
    for (const SplitSlice& sl : slices) {
        const SgbSplitSizes z = sgbk_split_sizes(sl.q1 - sl.q0, sl.rows, sl.maxlen, m.S);
        SgbPlanDev pd = pd0;
        pd.ostart += sl.q0; pd.len += sl.q0; pd.nseq = sl.q1 - sl.q0;
        SGB_CUDA(sgbk_split_decode(pd, sl.q1 - sl.q0, in, d_tabs, m, W.dev(z), d_state_out, sm_count, st,
                                   first ? ev_prep : nullptr, first ? ev_mid : nullptr));
        first = false;
    }
    return 0;
}

int upload_plan(Slot& S, cudaStream_t st, SgbPlanDev& pd)
{
    const HostPlan& hp = S.hp;
    size_t ns = (size_t)hp.nseq;
    size_t ng = hp.goff.size();
    if (S.goff.reserve(ng * sizeof(long long) + 16)) return 1;
    if (S.ostart.reserve(ns * sizeof(long long) + 16)) return 1;
    if (S.len.reserve(ns * sizeof(int) + 16)) return 1;
    if (ns) {
        const size_t b_goff = ng * sizeof(long long), b_ost = ns * sizeof(long long), b_len = ns * sizeof(int);
        if (S.pin.reserve(b_goff + b_ost + b_len)) return 1;
        char* h = (char*)S.pin.p;
        memcpy(h, hp.goff.data(), b_goff);
        memcpy(h + b_goff, hp.ostart.data(), b_ost);
        memcpy(h + b_goff + b_ost, hp.len.data(), b_len);
        SGB_CUDA(cudaMemcpyAsync(S.goff.p, h, b_goff, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(S.ostart.p, h + b_goff, b_ost, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(S.len.p, h + b_goff + b_ost, b_len, cudaMemcpyHostToDevice, st));
    }
    pd.goff = (const long long*)S.goff.p;
    pd.ostart = (const long long*)S.ostart.p;
    pd.len = (const int*)S.len.p;
    pd.nseq = hp.nseq;
    pd.ngroups = (int)((hp.nseq + 31) / 32);
    return 0;
}

long long chunk_bins_target()
{
    const char* e = getenv("SGB_CHUNK_BINS");
    if (e) {
        long long v = atoll(e);
        if (v > 0) return v;
    }
    return 8ll << 20;
}

// the caller's host columns, each with its own element type (SGB_T_*), and where the decoded states go
struct HostCols {
    const void* n; int n_type;
    const void* x; int x_type;
    const void* cn; int cn_type;
    void* out; int out_type;
};

bool host_ptr_is_pinned(const void* p)
{
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}

int host_threads_default(int share)
{
    const char* e = getenv("SGB_HOST_THREADS");
    int v = e ? atoi(e) : 0;
    if (v <= 0) {
        v = (int)std::thread::hardware_concurrency();
        if (v <= 0) v = 4;
        if (v > 16) v = 16;
    }
    v /= (share > 1 ? share : 1);
    return v < 1 ? 1 : v;
}

// Decode sequences [s_begin, s_end) of a host-resident batch on the current device.
//
// Chunks of the batch stream through two slots (copy / compute overlap).  Two ways of feeding a chunk:
//   direct  cudaMemcpyAsync straight from the caller's columns (int32 counts, int32 / uint8 states): right when they are
//           page-locked and host memory bandwidth is the scarce resource (several ranks on one node)
//   staged  worker threads copy the chunk into the slot's page-locked buffers, narrowing counts to uint16 (int32 if a
//           value does not fit) and states to uint8 -- 5 instead of 12-24 bytes per cell-bin over PCIe -- and convert
//           the decoded states into the caller's element type afterwards.  The only way for pageable memory (R vectors)
//           and for double / uint16 columns.  SGB_HOST_STAGE=always|never|auto (auto: staged unless direct is possible
//           AND the columns are page-locked; SGB_HOST_NARROW=1 stages page-locked columns too).  Measured on the 16-vCPU
//           B200 box (57 M cell-bins): page-locked int32 direct 13.6 ms, staged 18.1 ms -- the staging pass is bound by the
//           host's ~60 GB/s memory system (12 B read + 5 B written per cell-bin), slower than the DMA engine reading the
//           12 B itself -- pageable int32 staged 16.3 ms, pageable double in / double out 27.8 ms.
int decode_host_range(Ctx* C, const sgb_params* P, const SgbModelArgs& m, const HostCols& H, const int64_t* seg_off,
                      int64_t s_begin, int64_t s_end, double* loglik_out, int share)
{
    if (s_end <= s_begin) return 0;
    const bool direct_types = H.n_type == SGB_T_I32 && H.x_type == SGB_T_I32 && (H.cn_type == SGB_T_I32 || H.cn_type == SGB_T_U8) &&
                              H.out_type == SGB_T_U8;
    bool staged = true;
    {
        const char* e = getenv("SGB_HOST_STAGE");
        const char* nw = getenv("SGB_HOST_NARROW");
        const bool pinned = direct_types && host_ptr_is_pinned(H.n) && host_ptr_is_pinned(H.x) && host_ptr_is_pinned(H.cn) &&
                            host_ptr_is_pinned(H.out);
        if (e && !strcmp(e, "never")) staged = !direct_types;
        else if (e && !strcmp(e, "always")) staged = true;
        else staged = !pinned || (nw && !strcmp(nw, "1"));
    }
    if (loglik_out && !direct_types) return fail("loglik_out needs int32 count columns and uint8 output");
    // the `l` matrix kernel and the chained decode read int32 counts: no narrowing for them
    // (uint16 counts are read by k_emit only: a caller who forces the fused schedule gets int32 staging)
    const char* sched_env = getenv("SGB_SCHED");
    const bool narrow_ok = staged && !loglik_out && m.decode == SGB_DECODE_REFERENCE && !(sched_env && !strcmp(sched_env, "fused"));
    if (staged && !C->pool) C->pool.reset(new SgbPool(host_threads_default(share)));
    const int cn_itemsize = H.cn_type == SGB_T_I32 ? 4 : 1;

    SGB_CUDA(cudaMemsetAsync(C->d_err, 0, sizeof(int), C->main));
    SGB_CUDA(sgbk_setup(C->d_tabs, m, P->eps, P->Abias, P->selftransitionprob, C->main));
    SGB_CUDA(cudaEventRecord(C->ev_setup, C->main));
    const long long target_full = chunk_bins_target();
    const bool explicit_chunk = getenv("SGB_CHUNK_BINS") != nullptr;      // tests pin the chunk size exactly
    const int S = m.S;
    int host_bits = 0;
    C->slot[0].pend_rows = 0; C->slot[1].pend_rows = 0;
    // copies the decoded states a slot holds in its staging buffer into the caller's array (all workers)
    auto unstage = [&](Slot& SL) {
        if (SL.pend_rows <= 0) return;
        const int64_t rows = SL.pend_rows, r0 = SL.pend_r0;
        const uint8_t* src = (const uint8_t*)SL.stage_out.p;
        C->pool->run([&](int w, int nw) {
            const int64_t a = rows * w / nw, b = rows * (w + 1) / nw;
            if (b > a) sgb_unstage_states(src + a, b - a, H.out, H.out_type, r0 + a);
        });
        SL.pend_rows = 0;
    };
    int k = 0;
    int64_t s0 = s_begin;
    const int64_t row_end = seg_off[s_end];
    while (s0 < s_end) {
        int64_t s1 = s0;
        long long bins = 0;
        // a chunk should (a) be big enough to amortise launches and (b) hold enough sequences to give every
        // resident warp of the device a group (long sequences need more bins for that), within a memory cap;
        // the first chunks ramp up (1/8, 1/4, 1/2 of the target) and the last ones ramp down so that little of the
        // pipeline's fill and drain time is exposed
        long long target = target_full;
        if (!explicit_chunk) {
            if (k < 3) target = target_full >> (3 - k);
            const long long remaining = (long long)(row_end - seg_off[s0]);
            if (remaining < 2 * target) target = std::max<long long>(remaining / 2, target_full >> 3);
        }
        const long long seq_target = explicit_chunk ? 0 : (k < 3 ? 0 : (long long)C->sm_count * 8 * 32);
        const long long bins_cap = explicit_chunk ? target : (128ll << 20);
        while (s1 < s_end) {
            const long long L = seg_off[s1 + 1] - seg_off[s1];
            if (bins > 0 && bins + L > bins_cap) break;
            if (bins > 0 && bins + L > target && (s1 - s0) >= seq_target) break;
            bins += L;
            s1++;
        }
        Slot& SL = C->slot[k & 1];
        cudaStream_t st = SL.st;
        SGB_CUDA(cudaStreamSynchronize(st));          // slot buffers (and its host plan) are free again
        const int64_t r0 = seg_off[s0], r1 = seg_off[s1];
        const long long rows = (long long)(r1 - r0);
        const int use_split0 = want_split(m, s1 - s0);
        if (build_host_plan(seg_off, s0, s1, r0, SL.hp, (long long)C->sm_count * 16,
                            C->group_hint ? C->group_hint : ((use_split0 || narrow_ok) ? 32 : 0)))
            return 1;
        if (rows > 0) {
            if (SL.n.reserve((size_t)rows * 4) || SL.x.reserve((size_t)rows * 4) ||
                SL.cn.reserve((size_t)rows * cn_itemsize) || SL.out.reserve((size_t)rows))
                return 1;
            if (m.decode == SGB_DECODE_CHAINED &&
                (SL.bp.reserve((size_t)rows * sgb_bp_stride(S) + 16) || SL.last_state.reserve((size_t)SL.hp.nseq * 4 + 16)))
                return 1;
            if (loglik_out && SL.loglik.reserve((size_t)rows * S * 8)) return 1;
            SgbPlanDev pd;
            if (upload_plan(SL, st, pd)) return 1;
            int kind = cn_itemsize == 4 ? SGB_IN_I32_I32 : SGB_IN_I32_U8;
            int dev_cn_itemsize = cn_itemsize;
            if (staged) {
                const size_t off_x = ((size_t)rows * 4 + 255) & ~(size_t)255, off_c = 2 * off_x;
                // growing stage_out would drop the states of chunk k-2 still waiting in it: hand those over first
                if (SL.pend_rows > 0 && (size_t)rows + 256 > SL.stage_out.cap) unstage(SL);
                if (SL.stage_in.reserve(off_c + (size_t)rows + 256) || SL.stage_out.reserve((size_t)rows + 256)) return 1;
                char* base = (char*)SL.stage_in.p;
                std::atomic<int> bits(0);
                bool narrow = narrow_ok;
                // one pass of all workers: the states waiting in this slot go out to the caller, this chunk comes in
                const int64_t prows = SL.pend_rows, pr0 = SL.pend_r0;
                const uint8_t* psrc = (const uint8_t*)SL.stage_out.p;
                C->pool->run([&](int w, int nw) {
                    if (prows > 0) {
                        const int64_t a = prows * w / nw, b = prows * (w + 1) / nw;
                        if (b > a) sgb_unstage_states(psrc + a, b - a, H.out, H.out_type, pr0 + a);
                    }
                    const int64_t a = r0 + rows * w / nw, b = r0 + rows * (w + 1) / nw;
                    if (b <= a) return;
                    int bb = 0;
                    if (narrow) {
                        bb |= sgb_stage_count_u16(H.n, H.n_type, a, b, (uint16_t*)base + (a - r0));
                        bb |= sgb_stage_count_u16(H.x, H.x_type, a, b, (uint16_t*)(base + off_x) + (a - r0));
                    } else {
                        bb |= sgb_stage_count_i32(H.n, H.n_type, a, b, (int32_t*)base + (a - r0));
                        bb |= sgb_stage_count_i32(H.x, H.x_type, a, b, (int32_t*)(base + off_x) + (a - r0));
                    }
                    bb |= sgb_stage_cn_u8(H.cn, H.cn_type, a, b, (uint8_t*)(base + off_c) + (a - r0));
                    if (bb) bits.fetch_or(bb);
                });
                SL.pend_rows = 0;
                if (narrow && (bits.load() & SGB_STG_WIDE)) {
                    // a count beyond uint16 (or negative): this chunk goes over as int32 and the device validates it
                    narrow = false;
                    C->pool->run([&](int w, int nw) {
                        const int64_t a = r0 + rows * w / nw, b = r0 + rows * (w + 1) / nw;
                        if (b <= a) return;
                        const int bb = sgb_stage_count_i32(H.n, H.n_type, a, b, (int32_t*)base + (a - r0)) |
                                       sgb_stage_count_i32(H.x, H.x_type, a, b, (int32_t*)(base + off_x) + (a - r0));
                        if (bb) bits.fetch_or(bb);
                    });
                }
                host_bits |= bits.load() & (SGB_STG_NOT_INT | SGB_STG_BAD_CN);
                const size_t esz = narrow ? 2 : 4;
                SGB_CUDA(cudaMemcpyAsync(SL.n.p, base, (size_t)rows * esz, cudaMemcpyHostToDevice, st));
                SGB_CUDA(cudaMemcpyAsync(SL.x.p, base + off_x, (size_t)rows * esz, cudaMemcpyHostToDevice, st));
                SGB_CUDA(cudaMemcpyAsync(SL.cn.p, base + off_c, (size_t)rows, cudaMemcpyHostToDevice, st));
                kind = narrow ? SGB_IN_U16_U8 : SGB_IN_I32_U8;
                dev_cn_itemsize = 1;
            } else {
                SGB_CUDA(cudaMemcpyAsync(SL.n.p, (const int32_t*)H.n + r0, (size_t)rows * 4, cudaMemcpyHostToDevice, st));
                SGB_CUDA(cudaMemcpyAsync(SL.x.p, (const int32_t*)H.x + r0, (size_t)rows * 4, cudaMemcpyHostToDevice, st));
                SGB_CUDA(cudaMemcpyAsync(SL.cn.p, (const char*)H.cn + (size_t)r0 * cn_itemsize, (size_t)rows * cn_itemsize,
                                         cudaMemcpyHostToDevice, st));
            }
            SGB_CUDA(cudaStreamWaitEvent(st, C->ev_setup, 0));
            const int use_split = (m.decode == SGB_DECODE_REFERENCE) && (use_split0 || kind == SGB_IN_U16_U8);
            if (use_split) {
                const SgbColsIn in = {SL.n.p, SL.x.p, SL.cn.p, kind, rows, C->d_err};
                std::vector<SplitSlice> slices;
                if (run_split(SL.split, slices, SL.hp, pd, in, C->d_tabs, m, (uint8_t*)SL.out.p, C->sm_count, st, nullptr, nullptr))
                    return 1;
            } else {
                SgbRawIn raw = {(const int32_t*)SL.n.p, (const int32_t*)SL.x.p, SL.cn.p, dev_cn_itemsize, rows, C->d_err};
                if (dev_cn_itemsize == 4) {
                    // an int32 cn column (R integers) is narrowed to uint8 on the device when that lets k_decode read in place
                    SgbRawIn probe = raw;
                    probe.cn = nullptr; probe.cn_itemsize = 1;
                    if (sgbk_resolve_stage(m, C->stage, probe) == 2) {
                        if (SL.cn8.reserve((size_t)rows + 16)) return 1;
                        SGB_CUDA(sgbk_narrow_cn((const int32_t*)SL.cn.p, rows, (uint8_t*)SL.cn8.p, C->d_err, st));
                        raw.cn = SL.cn8.p; raw.cn_itemsize = 1;
                    }
                }
                const int stage = sgbk_resolve_stage(m, C->stage, raw);
                if (stage != 2 && SL.packed.reserve((size_t)SL.hp.packed_len * 8 + 16)) return 1;    // packed records: only off the RAW path
                if (stage != 2)
                    SGB_CUDA(sgbk_pack((const int32_t*)SL.n.p, (const int32_t*)SL.x.p, SL.cn.p, dev_cn_itemsize, pd,
                                       (unsigned long long*)SL.packed.p, C->d_err, st));
                SGB_CUDA(sgbk_decode(pd, (const unsigned long long*)SL.packed.p, raw, C->d_tabs, m, (uint8_t*)SL.out.p,
                                     (uint8_t*)SL.bp.p, (int*)SL.last_state.p, SL.counter, C->sm_count, stage, st));
                if (m.decode == SGB_DECODE_CHAINED)
                    SGB_CUDA(sgbk_backtrace(pd, (const uint8_t*)SL.bp.p, (const int*)SL.last_state.p, S, (uint8_t*)SL.out.p, st));
            }
            if (staged) {
                SGB_CUDA(cudaMemcpyAsync(SL.stage_out.p, SL.out.p, (size_t)rows, cudaMemcpyDeviceToHost, st));
                SL.pend_r0 = r0; SL.pend_rows = rows;
            } else {
                SGB_CUDA(cudaMemcpyAsync((uint8_t*)H.out + r0, SL.out.p, (size_t)rows, cudaMemcpyDeviceToHost, st));
            }
            if (loglik_out) {
                SGB_CUDA(sgbk_loglik((const int32_t*)SL.n.p, (const int32_t*)SL.x.p, SL.cn.p, dev_cn_itemsize, pd,
                                     C->d_tabs, m, (double*)SL.loglik.p, st));
                SGB_CUDA(cudaMemcpyAsync(loglik_out + (size_t)r0 * S, SL.loglik.p, (size_t)rows * S * 8,
                                         cudaMemcpyDeviceToHost, st));
            }
        }
        s0 = s1;
        k++;
    }
    // drain in slot order: the older slot's states go out while the younger one still computes
    for (int j = 0; j < 2; j++) {
        Slot& SL = C->slot[(k + j) & 1];
        SGB_CUDA(cudaStreamSynchronize(SL.st));
        if (staged) unstage(SL);
    }
    int bits = 0;
    SGB_CUDA(cudaMemcpy(&bits, C->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (host_bits & SGB_STG_NOT_INT) bits |= SGB_ERR_NOT_INTEGER;
    if (host_bits & SGB_STG_BAD_CN) bits |= SGB_ERR_BAD_CN;
    if (bits) return fail(err_bits_text(bits));
    return 0;
}

}  // namespace

void sgb_internal_set_error(const char* msg) { t_err = msg ? msg : ""; }

// ====================================================================== C ABI
extern "C" {

const char* sgb_last_error(void) { return t_err.c_str(); }

int sgb_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        fail(std::string("CUDA unusable: ") + cudaGetErrorString(e));
        return 0;
    }
    return n;
}

int sgb_init(int device)
{
    if (device >= 0) SGB_CUDA(cudaSetDevice(device));
    Ctx* C;
    return get_ctx(&C);
}

int sgb_shutdown(void)
{
    std::lock_guard<std::mutex> lk(g_mu);
    int cur = 0;
    cudaGetDevice(&cur);
    for (int d = 0; d < 64; d++) {
        Ctx* c = g_ctx[d];
        if (!c) continue;
        cudaSetDevice(d);
        for (int i = 0; i < 2; i++) {
            Slot& s = c->slot[i];
            if (s.st) { cudaStreamSynchronize(s.st); cudaStreamDestroy(s.st); }
            s.n.release(); s.x.release(); s.cn.release(); s.cn8.release(); s.out.release(); s.packed.release(); s.bp.release();
            s.last_state.release(); s.pin.release(); s.goff.release(); s.ostart.release(); s.len.release(); s.loglik.release();
            s.split.release(); s.stage_in.release(); s.stage_out.release();
            if (s.counter) cudaFree(s.counter);
        }
        c->scratch_a.release(); c->scratch_b.release(); c->scratch_c.release(); c->scratch_d.release();
        if (c->main) cudaStreamDestroy(c->main);
        if (c->ev_setup) cudaEventDestroy(c->ev_setup);
        if (c->d_tabs) cudaFree(c->d_tabs);
        if (c->d_err) cudaFree(c->d_err);
        delete c;
        g_ctx[d] = nullptr;
    }
    cudaSetDevice(cur);
    return 0;
}

static int decode_typed(const sgb_params* params, int64_t N, const HostCols& H, int64_t nseg, const int64_t* seg_off,
                        double* loglik_out, int32_t n_devices)
{
    SgbModelArgs m;
    if (check_params(params, m)) return 1;
    if (N < 0 || nseg < 0) return fail("negative size");
    if (!sgb_type_size(H.n_type) || !sgb_type_size(H.x_type) || !sgb_type_size(H.cn_type))
        return fail("column element type must be one of SGB_T_U8, SGB_T_U16, SGB_T_I32, SGB_T_F64");
    if (!sgb_type_size(H.out_type)) return fail("state_out element type must be one of SGB_T_U8, SGB_T_U16, SGB_T_I32, SGB_T_F64");
    if (nseg > 0 && !seg_off) return fail("seg_off is NULL");
    if (N > 0 && (!H.n || !H.x || !H.cn || !H.out)) return fail("NULL data pointer");
    if (nseg > 0) {
        if (seg_off[0] < 0 || seg_off[nseg] > N) return fail("seg_off out of range");
        for (int64_t s = 0; s < nseg; s++)
            if (seg_off[s + 1] < seg_off[s]) return fail("seg_off must be non-decreasing");
    }
    if (n_devices <= 1) {
        Ctx* C;
        if (get_ctx(&C)) return 1;
        return decode_host_range(C, params, m, H, seg_off, 0, nseg, loglik_out, 1);
    }
    if (loglik_out) return fail("loglik_out is single-device only");
    int avail = sgb_device_count();
    if (n_devices > avail) return fail("n_devices must be between 1 and the number of visible CUDA devices");
    if (nseg == 0) return 0;
    // contiguous blocks of sequences, balanced by bins
    std::vector<int64_t> cut((size_t)n_devices + 1, nseg);
    cut[0] = 0;
    const int64_t total = seg_off[nseg] - seg_off[0];
    int64_t s = 0;
    for (int d = 1; d < n_devices; d++) {
        const int64_t want = seg_off[0] + total * d / n_devices;
        while (s < nseg && seg_off[s] < want) s++;
        cut[(size_t)d] = s;
    }
    int prev = 0;
    cudaGetDevice(&prev);
    std::vector<int> rc((size_t)n_devices, 0);
    std::vector<std::string> msg((size_t)n_devices);
    std::vector<std::thread> th;
    for (int d = 0; d < n_devices; d++) {
        th.emplace_back([&, d]() {
            if (cudaSetDevice(d) != cudaSuccess) { rc[(size_t)d] = 1; msg[(size_t)d] = "cudaSetDevice failed"; return; }
            Ctx* C;
            if (get_ctx(&C)) { rc[(size_t)d] = 1; msg[(size_t)d] = t_err; return; }
            rc[(size_t)d] = decode_host_range(C, params, m, H, seg_off, cut[(size_t)d], cut[(size_t)d + 1], nullptr, n_devices);
            if (rc[(size_t)d]) msg[(size_t)d] = t_err;
        });
    }
    for (auto& t : th) t.join();
    cudaSetDevice(prev);
    for (int d = 0; d < n_devices; d++)
        if (rc[(size_t)d]) return fail("device " + std::to_string(d) + ": " + msg[(size_t)d]);
    return 0;
}

int sgb_hmm_decode(const sgb_params* params, int64_t N, const int32_t* n, const int32_t* x, const void* cn,
                   int32_t cn_itemsize, int64_t nseg, const int64_t* seg_off, uint8_t* state_out,
                   double* loglik_out)
{
    if (cn_itemsize != 1 && cn_itemsize != 4) return fail("cn_itemsize must be 1 (uint8) or 4 (int32)");
    const HostCols H = {n, SGB_T_I32, x, SGB_T_I32, cn, cn_itemsize == 4 ? SGB_T_I32 : SGB_T_U8, state_out, SGB_T_U8};
    return decode_typed(params, N, H, nseg, seg_off, loglik_out, 1);
}

int sgb_hmm_decode_multi(const sgb_params* params, int32_t n_devices, int64_t N, const int32_t* n,
                         const int32_t* x, const void* cn, int32_t cn_itemsize, int64_t nseg,
                         const int64_t* seg_off, uint8_t* state_out)
{
    if (cn_itemsize != 1 && cn_itemsize != 4) return fail("cn_itemsize must be 1 (uint8) or 4 (int32)");
    if (n_devices < 1) return fail("n_devices must be between 1 and the number of visible CUDA devices");
    const HostCols H = {n, SGB_T_I32, x, SGB_T_I32, cn, cn_itemsize == 4 ? SGB_T_I32 : SGB_T_U8, state_out, SGB_T_U8};
    return decode_typed(params, N, H, nseg, seg_off, nullptr, n_devices);
}

int sgb_hmm_decode_typed(const sgb_params* params, int64_t N, const void* n, int32_t n_type, const void* x, int32_t x_type,
                         const void* cn, int32_t cn_type, int64_t nseg, const int64_t* seg_off, void* state_out,
                         int32_t out_type, int32_t n_devices)
{
    if (n_devices < 1) return fail("n_devices must be between 1 and the number of visible CUDA devices");
    const HostCols H = {n, n_type, x, x_type, cn, cn_type, state_out, out_type};
    return decode_typed(params, N, H, nseg, seg_off, nullptr, n_devices);
}

// ---------------------------------------------------------------- device-resident path
struct sgb_plan {
    int device = 0;
    int64_t N = 0;
    int n_states = 0, decode = 0;
    int64_t nseg = 0;
    HostPlan hp;
    DevBuf goff, ostart, len, packed, cn8, bp, last_state;
    SplitWork split;                        // split schedule scratch (reserved by the first call that takes it)
    std::vector<SplitSlice> slices;
    unsigned int* counter = nullptr;
    int* d_err = nullptr;
    SgbDevTables* d_tabs = nullptr;
    SgbPlanDev pd;
    // optional per-call kernel timing: ring of event quadruples (start, after pack / split preparation, after k_emit,
    // after k_decode / k_recur)
    bool timing = false;
    static const int kRing = 64;
    cudaEvent_t ev[kRing][4];
    bool ev_ready = false;
    long long calls = 0;
    int last_stage = -1;
};

int sgb_plan_create(int64_t N, int64_t nseg, const int64_t* seg_off, int32_t n_states, int32_t decode,
                    sgb_plan** plan_out)
{
    if (!plan_out) return fail("plan_out is NULL");
    *plan_out = nullptr;
    if (N < 0 || nseg < 0 || (nseg > 0 && !seg_off)) return fail("bad plan arguments");
    if (n_states < 1 || n_states > SGB_MAX_STATES) return fail("n_states (maxCN+1) must be in 1..32");
    if (nseg > 0 && (seg_off[0] < 0 || seg_off[nseg] > N)) return fail("seg_off out of range");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    sgb_plan* pl = new sgb_plan();
    pl->device = C->device; pl->N = N; pl->n_states = n_states; pl->decode = decode; pl->nseg = nseg;
    {
        // the split schedule (the default for decode = reference) wants full groups of 32: its emission kernel is not
        // starved by small batches, and narrow groups would only add empty plan entries
        SgbModelArgs probe; probe.decode = decode;
        const int hint = C->group_hint ? C->group_hint : (want_split(probe, nseg) ? 32 : 0);
        if (build_host_plan(seg_off, 0, nseg, 0, pl->hp, (long long)C->sm_count * 16, hint)) { delete pl; return 1; }
    }
    size_t ns = (size_t)pl->hp.nseq;
    size_t ng = pl->hp.goff.size();
    // the narrowed-cn buffer (1 B per cell-bin) is always reserved; the packed-record buffer (8 B per cell-bin) is only
    // needed off the in-place path: reserved now when that is already certain, else by sgb_plan_reserve_packed or by the
    // first decode call that turns out to need it
    int bad = pl->goff.reserve(ng * 8 + 16) || pl->ostart.reserve(ns * 8 + 16) || pl->len.reserve(ns * 4 + 16) ||
              pl->cn8.reserve((size_t)N + 16) ||
              ((C->stage != 2 || decode == SGB_DECODE_CHAINED) && pl->packed.reserve((size_t)pl->hp.packed_len * 8 + 16));
    if (!bad && decode == SGB_DECODE_CHAINED)
        bad = pl->bp.reserve((size_t)N * sgb_bp_stride(n_states) + 16) || pl->last_state.reserve(ns * 4 + 16);
    if (bad) { sgb_plan_destroy(pl); return 1; }
    cudaError_t e = cudaSuccess;
    if (ns) {
        e = cudaMemcpy(pl->goff.p, pl->hp.goff.data(), ng * 8, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(pl->ostart.p, pl->hp.ostart.data(), ns * 8, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(pl->len.p, pl->hp.len.data(), ns * 4, cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaMalloc(&pl->counter, sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_err, sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(pl->d_err, 0, sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&pl->d_tabs, sizeof(SgbDevTables));
    if (e != cudaSuccess) {
        sgb_plan_destroy(pl);
        return fail(std::string("plan allocation failed: ") + cudaGetErrorString(e));
    }
    pl->pd.goff = (const long long*)pl->goff.p;
    pl->pd.ostart = (const long long*)pl->ostart.p;
    pl->pd.len = (const int*)pl->len.p;
    pl->pd.nseq = pl->hp.nseq;
    pl->pd.ngroups = (int)((pl->hp.nseq + 31) / 32);
    *plan_out = pl;
    return 0;
}

int sgb_plan_destroy(sgb_plan* pl)
{
    if (!pl) return 0;
    if (pl->ev_ready)
        for (int i = 0; i < sgb_plan::kRing; i++)
            for (int j = 0; j < 4; j++) cudaEventDestroy(pl->ev[i][j]);
    pl->split.release();
    pl->goff.release(); pl->ostart.release(); pl->len.release(); pl->packed.release(); pl->cn8.release(); pl->bp.release();
    pl->last_state.release();
    if (pl->counter) cudaFree(pl->counter);
    if (pl->d_err) cudaFree(pl->d_err);
    if (pl->d_tabs) cudaFree(pl->d_tabs);
    delete pl;
    return 0;
}

int sgb_plan_reserve_packed(sgb_plan* pl)
{
    if (!pl) return fail("plan is NULL");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    if (C->device != pl->device) return fail("plan belongs to another device");
    return pl->packed.reserve((size_t)pl->hp.packed_len * 8 + 16);
}

int64_t sgb_plan_scratch_bytes(const sgb_plan* pl)
{
    if (!pl) return 0;
    return (int64_t)(pl->split.bytes() + pl->packed.cap + pl->cn8.cap + pl->bp.cap + pl->last_state.cap + pl->goff.cap + pl->ostart.cap + pl->len.cap);
}

int sgb_hmm_decode_device(const sgb_params* params, sgb_plan* pl, const int32_t* d_n, const int32_t* d_x,
                          const void* d_cn, int32_t cn_itemsize, uint8_t* d_state_out, double* d_loglik_out,
                          void* stream)
{
    SgbModelArgs m;
    if (check_params(params, m)) return 1;
    if (!pl) return fail("plan is NULL");
    if (cn_itemsize != 1 && cn_itemsize != 4) return fail("cn_itemsize must be 1 (uint8) or 4 (int32)");
    if (m.S != pl->n_states) return fail("params.n_states differs from the plan's");
    if (m.decode == SGB_DECODE_CHAINED && pl->decode != SGB_DECODE_CHAINED)
        return fail("plan was created without backpointer scratch (decode = chained)");
    if (pl->hp.nseq == 0) return 0;
    if (!d_n || !d_x || !d_cn || !d_state_out) return fail("NULL device pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    if (C->device != pl->device) return fail("plan belongs to another device");
    cudaStream_t st = (cudaStream_t)stream;
    SGB_CUDA(sgbk_setup(pl->d_tabs, m, params->eps, params->Abias, params->selftransitionprob, st));
    const int slot = (int)(pl->calls % sgb_plan::kRing);
    if (pl->timing) SGB_CUDA(cudaEventRecord(pl->ev[slot][0], st));
    if (want_split(m, pl->nseg)) {
        // ---- split schedule: k_emit (lanes over bins) + k_recur (lane per sequence), sliced if the batch is huge
        const SgbColsIn in = {d_n, d_x, d_cn, cn_itemsize == 4 ? SGB_IN_I32_I32 : SGB_IN_I32_U8, (long long)pl->N, pl->d_err};
        if (run_split(pl->split, pl->slices, pl->hp, pl->pd, in, pl->d_tabs, m, d_state_out, C->sm_count, st,
                      pl->timing ? pl->ev[slot][1] : nullptr, pl->timing ? pl->ev[slot][2] : nullptr))
            return 1;
        pl->last_stage = 4;
        if (pl->timing) { SGB_CUDA(cudaEventRecord(pl->ev[slot][3], st)); pl->calls++; }
        if (d_loglik_out)
            SGB_CUDA(sgbk_loglik(d_n, d_x, d_cn, cn_itemsize, pl->pd, pl->d_tabs, m, d_loglik_out, st));
        return 0;
    }
    SgbRawIn raw = {d_n, d_x, d_cn, cn_itemsize, (long long)pl->N, pl->d_err};
    bool narrowed = false;
    if (cn_itemsize == 4 && ((unsigned long long)d_cn & 15ull) == 0) {
        SgbRawIn probe = raw;
        probe.cn = nullptr; probe.cn_itemsize = 1;
        if (sgbk_resolve_stage(m, C->stage, probe) == 2) {
            SGB_CUDA(sgbk_narrow_cn((const int32_t*)d_cn, (long long)pl->N, (uint8_t*)pl->cn8.p, pl->d_err, st));
            raw.cn = pl->cn8.p; raw.cn_itemsize = 1;
            narrowed = true;
        }
    }
    const int stage = sgbk_resolve_stage(m, C->stage, raw);
    if (stage != 2 && pl->packed.reserve((size_t)pl->hp.packed_len * 8 + 16)) return 1;
    pl->last_stage = (stage == 2 && narrowed) ? 3 : stage;
    if (stage != 2)
        SGB_CUDA(sgbk_pack(d_n, d_x, d_cn, cn_itemsize, pl->pd, (unsigned long long*)pl->packed.p, pl->d_err, st));
    if (pl->timing) { SGB_CUDA(cudaEventRecord(pl->ev[slot][1], st)); SGB_CUDA(cudaEventRecord(pl->ev[slot][2], st)); }
    SGB_CUDA(sgbk_decode(pl->pd, (const unsigned long long*)pl->packed.p, raw, pl->d_tabs, m, d_state_out,
                         (uint8_t*)pl->bp.p, (int*)pl->last_state.p, pl->counter, C->sm_count, stage, st));
    if (pl->timing) { SGB_CUDA(cudaEventRecord(pl->ev[slot][3], st)); pl->calls++; }
    if (m.decode == SGB_DECODE_CHAINED)
        SGB_CUDA(sgbk_backtrace(pl->pd, (const uint8_t*)pl->bp.p, (const int*)pl->last_state.p, m.S, d_state_out, st));
    if (d_loglik_out)
        SGB_CUDA(sgbk_loglik(d_n, d_x, d_cn, cn_itemsize, pl->pd, pl->d_tabs, m, d_loglik_out, st));
    return 0;
}

int sgb_plan_last_staging(const sgb_plan* pl) { return pl ? pl->last_stage : -1; }

int sgb_plan_enable_timing(sgb_plan* pl, int32_t on)
{
    if (!pl) return fail("plan is NULL");
    if (on && !pl->ev_ready) {
        for (int i = 0; i < sgb_plan::kRing; i++)
            for (int j = 0; j < 4; j++) SGB_CUDA(cudaEventCreate(&pl->ev[i][j]));
        pl->ev_ready = true;
    }
    pl->timing = on != 0;
    pl->calls = 0;
    return 0;
}

int sgb_plan_read_timing3(sgb_plan* pl, int32_t max_records, float* prep_ms, float* emit_ms, float* decode_ms, int32_t* n_out)
{
    if (!pl || !prep_ms || !emit_ms || !decode_ms || !n_out) return fail("NULL pointer");
    *n_out = 0;
    if (!pl->ev_ready) return 0;
    long long have = pl->calls < sgb_plan::kRing ? pl->calls : sgb_plan::kRing;
    if (have > max_records) have = max_records;
    for (long long i = 0; i < have; i++) {
        int slot = (int)((pl->calls - have + i) % sgb_plan::kRing);
        SGB_CUDA(cudaEventSynchronize(pl->ev[slot][3]));
        SGB_CUDA(cudaEventElapsedTime(&prep_ms[i], pl->ev[slot][0], pl->ev[slot][1]));
        SGB_CUDA(cudaEventElapsedTime(&emit_ms[i], pl->ev[slot][1], pl->ev[slot][2]));
        SGB_CUDA(cudaEventElapsedTime(&decode_ms[i], pl->ev[slot][2], pl->ev[slot][3]));
    }
    *n_out = (int32_t)have;
    return 0;
}

int sgb_plan_read_timing(sgb_plan* pl, int32_t max_records, float* pack_ms, float* decode_ms, int32_t* n_out)
{
    if (!pl || !pack_ms || !decode_ms || !n_out) return fail("NULL pointer");
    std::vector<float> em((size_t)(max_records > 0 ? max_records : 1));
    if (sgb_plan_read_timing3(pl, max_records, pack_ms, em.data(), decode_ms, n_out)) return 1;
    for (int i = 0; i < *n_out; i++) decode_ms[i] += em[(size_t)i];          // two-phase view: preparation | kernels
    return 0;
}

int sgb_plan_split_info(sgb_plan* pl, void* stream, int64_t* ngroups_out, int64_t* e_bytes_out)
{
    if (!pl || !ngroups_out || !e_bytes_out) return fail("NULL pointer");
    *ngroups_out = 0; *e_bytes_out = 0;
    if (pl->last_stage != 4 || pl->slices.empty() || !pl->split.misc.p) return 0;
    SGB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    // the device-side tables describe the LAST slice the plan decoded
    int ng = 0;
    SGB_CUDA(cudaMemcpy(&ng, pl->split.misc.p, sizeof(int), cudaMemcpyDeviceToHost));
    if (ng > 0) {
        long long off = 0;
        unsigned long long sz = 0;
        SGB_CUDA(cudaMemcpy(&off, (long long*)pl->split.g_eoff.p + (ng - 1), 8, cudaMemcpyDeviceToHost));
        SGB_CUDA(cudaMemcpy(&sz, (unsigned long long*)pl->split.g_size.p + (ng - 1), 8, cudaMemcpyDeviceToHost));
        *e_bytes_out = (int64_t)((unsigned long long)off + sz) * 32;
    }
    *ngroups_out = ng;
    return 0;
}

int sgb_plan_reserve_split(sgb_plan* pl)
{
    if (!pl) return fail("plan is NULL");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    if (C->device != pl->device) return fail("plan belongs to another device");
    if (pl->slices.empty()) pl->slices = split_slices(pl->hp.len, pl->n_states, split_budget_bytes());
    SgbSplitSizes zmax = sgbk_split_sizes(0, 0, 0, pl->n_states);
    for (const SplitSlice& sl : pl->slices) {
        const SgbSplitSizes z = sgbk_split_sizes(sl.q1 - sl.q0, sl.rows, sl.maxlen, pl->n_states);
        zmax.cls = std::max(zmax.cls, z.cls); zmax.perm = std::max(zmax.perm, z.perm); zmax.part = std::max(zmax.part, z.part);
        zmax.g_i32 = std::max(zmax.g_i32, z.g_i32); zmax.g_i64 = std::max(zmax.g_i64, z.g_i64);
        zmax.e_bytes = std::max(zmax.e_bytes, z.e_bytes);
    }
    return pl->split.reserve(zmax);
}

int sgb_plan_check(sgb_plan* pl, void* stream)
{
    if (!pl) return fail("plan is NULL");
    SGB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    int bits = 0;
    SGB_CUDA(cudaMemcpy(&bits, pl->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (bits) {
        cudaMemset(pl->d_err, 0, sizeof(int));
        return fail(err_bits_text(bits));
    }
    return 0;
}

// ---------------------------------------------------------------- resident columns (one upload, several passes)
// callHaplotypeSpecificCN decodes the same table twice (R/callHSCN.R:883-935: pass 1 at the default loherror, pass 2 at the
// inferred one, possibly beta-binomial), callAlleleSpecificCN once more on top: the columns are the same every time.
// sgb_cols keeps them on the device -- int32 counts, uint8 copy number, converted and validated on the way up by the staging
// workers -- so that every pass after the first costs the kernels and the copy of the states back.
struct sgb_cols {
    int device = 0;
    int64_t N = 0, nseg = 0;
    std::vector<int64_t> seg_off;
    DevBuf n, x, cn, out;
    sgb_plan* plan = nullptr;   // made by the first decode (its scratch depends on n_states and the decode rule)
};

int sgb_cols_free(sgb_cols* c)
{
    if (!c) return 0;
    if (c->plan) sgb_plan_destroy(c->plan);
    c->n.release(); c->x.release(); c->cn.release(); c->out.release();
    delete c;
    return 0;
}

int sgb_cols_upload(int64_t N, const void* n, int32_t n_type, const void* x, int32_t x_type, const void* cn, int32_t cn_type,
                    int64_t nseg, const int64_t* seg_off, sgb_cols** cols_out)
{
    if (!cols_out) return fail("cols_out is NULL");
    *cols_out = nullptr;
    if (N < 0 || nseg < 0 || (nseg > 0 && !seg_off)) return fail("bad arguments");
    if (N > 0 && (!n || !x || !cn)) return fail("NULL column");
    if (!sgb_type_size(n_type) || !sgb_type_size(x_type) || !sgb_type_size(cn_type)) return fail("unknown element type code");
    for (int64_t s = 0; s < nseg; s++)
        if (seg_off[s] > seg_off[s + 1]) return fail("seg_off must be non-decreasing");
    if (nseg > 0 && (seg_off[0] < 0 || seg_off[nseg] > N)) return fail("seg_off out of range");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    sgb_cols* c = new sgb_cols();
    c->device = C->device; c->N = N; c->nseg = nseg;
    c->seg_off.assign(seg_off, seg_off + (nseg > 0 ? nseg + 1 : 0));
    if (c->n.reserve((size_t)N * 4 + 16) || c->x.reserve((size_t)N * 4 + 16) || c->cn.reserve((size_t)N + 16) ||
        c->out.reserve((size_t)N + 16)) { sgb_cols_free(c); return 1; }
    // chunks of rows: workers convert into a page-locked slot, the slot is copied up while they fill the other one
    // (the page-locked slots are the context's chunk-staging buffers: allocated once per device -- page-locking 72 MB costs
    // ~40 ms, more than the upload itself -- and idle between calls, the library not being re-entrant)
    const int64_t CH = 4 << 20;
    PinBuf* slot[2] = {&C->slot[0].stage_in, &C->slot[1].stage_in};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    cudaStream_t st = nullptr;
    int rc = 0, host_bits = 0;
    if (N > 0) {
        if (!C->pool) C->pool.reset(new SgbPool(host_threads_default(1)));
        if (slot[0]->reserve((size_t)CH * 9) || slot[1]->reserve((size_t)CH * 9)) rc = 1;
        if (!rc && (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess ||
                    cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming) != cudaSuccess ||
                    cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming) != cudaSuccess))
            rc = fail("could not create the upload stream");
    }
    int k = 0;
    for (int64_t r0 = 0; !rc && r0 < N; r0 += CH, k++) {
        const int64_t rows = std::min<int64_t>(CH, N - r0);
        char* base = (char*)slot[k & 1]->p;
        if (k >= 2 && cudaEventSynchronize(ev[k & 1]) != cudaSuccess) { rc = fail("upload failed"); break; }
        std::atomic<int> bits(0);
        C->pool->run([&](int w, int nw) {
            const int64_t per = (rows + nw - 1) / nw;
            const int64_t a = r0 + std::min<int64_t>(rows, per * w), b = r0 + std::min<int64_t>(rows, per * (w + 1));
            if (a >= b) return;
            int bb = sgb_stage_count_i32(n, n_type, a, b, (int32_t*)base + (a - r0));
            bb |= sgb_stage_count_i32(x, x_type, a, b, (int32_t*)(base + (size_t)CH * 4) + (a - r0));
            bb |= sgb_stage_cn_u8(cn, cn_type, a, b, (uint8_t*)(base + (size_t)CH * 8) + (a - r0));
            bits |= bb;
        });
        host_bits |= bits.load();
        cudaError_t e = cudaMemcpyAsync((int32_t*)c->n.p + r0, base, (size_t)rows * 4, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync((int32_t*)c->x.p + r0, base + (size_t)CH * 4, (size_t)rows * 4, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync((uint8_t*)c->cn.p + r0, base + (size_t)CH * 8, (size_t)rows, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaEventRecord(ev[k & 1], st);
        if (e != cudaSuccess) rc = fail(std::string("upload failed: ") + cudaGetErrorString(e));
    }
    if (st) {
        if (cudaStreamSynchronize(st) != cudaSuccess && !rc) rc = fail("upload failed");
        cudaStreamDestroy(st);
    }
    for (int j = 0; j < 2; j++) if (ev[j]) cudaEventDestroy(ev[j]);
    if (!rc && (host_bits & (SGB_STG_NOT_INT | SGB_STG_BAD_CN))) {
        int eb = 0;
        if (host_bits & SGB_STG_NOT_INT) eb |= SGB_ERR_NOT_INTEGER;
        if (host_bits & SGB_STG_BAD_CN) eb |= SGB_ERR_BAD_CN;
        rc = fail(err_bits_text(eb));
    }
    if (rc) { sgb_cols_free(c); return 1; }
    *cols_out = c;
    return 0;
}

int sgb_cols_decode(sgb_cols* c, const sgb_params* params, void* state_out, int32_t out_type)
{
    if (!c) return fail("cols is NULL");
    SgbModelArgs m;
    if (check_params(params, m)) return 1;
    if (!sgb_type_size(out_type)) return fail("unknown element type code");
    if (c->N == 0) return 0;
    if (!state_out) return fail("state_out is NULL");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    if (C->device != c->device) return fail("columns live on another device");
    if (c->plan && (c->plan->n_states != m.S || (m.decode == SGB_DECODE_CHAINED && c->plan->decode != SGB_DECODE_CHAINED))) {
        sgb_plan_destroy(c->plan);
        c->plan = nullptr;
    }
    if (!c->plan && sgb_plan_create(c->N, c->nseg, c->seg_off.data(), m.S, m.decode, &c->plan)) return 1;
    cudaStream_t st = C->main;
    // bins that belong to no sequence keep state 0, as in sgb_hmm_decode
    SGB_CUDA(cudaMemsetAsync(c->out.p, 0, (size_t)c->N, st));
    if (sgb_hmm_decode_device(params, c->plan, (const int32_t*)c->n.p, (const int32_t*)c->x.p, c->cn.p, 1, (uint8_t*)c->out.p,
                              nullptr, (void*)st))
        return 1;
    if (sgb_plan_check(c->plan, (void*)st)) return 1;
    if (out_type == SGB_T_U8) {
        SGB_CUDA(cudaMemcpy(state_out, c->out.p, (size_t)c->N, cudaMemcpyDeviceToHost));
        return 0;
    }
    // other element types: states come down in page-locked chunks, the workers widen one chunk while the next one arrives
    const int64_t CH = 8 << 20;
    PinBuf& pin_out = C->slot[0].stage_out;       // (the context's page-locked output staging, idle between calls)
    if (pin_out.reserve((size_t)std::min<int64_t>(c->N, 2 * CH))) return 1;
    if (!C->pool) C->pool.reset(new SgbPool(host_threads_default(1)));
    const int64_t nch = (c->N + CH - 1) / CH;
    auto fetch = [&](int64_t k) {
        const int64_t r0 = k * CH, rows = std::min<int64_t>(CH, c->N - r0);
        return cudaMemcpyAsync((uint8_t*)pin_out.p + (k & 1) * CH, (const uint8_t*)c->out.p + r0, (size_t)rows, cudaMemcpyDeviceToHost, st);
    };
    SGB_CUDA(fetch(0));
    for (int64_t k = 0; k < nch; k++) {
        SGB_CUDA(cudaStreamSynchronize(st));
        if (k + 1 < nch) SGB_CUDA(fetch(k + 1));
        const int64_t r0 = k * CH, rows = std::min<int64_t>(CH, c->N - r0);
        const uint8_t* src = (const uint8_t*)pin_out.p + (k & 1) * CH;
        C->pool->run([&](int w, int nw) {
            const int64_t per = (rows + nw - 1) / nw;
            const int64_t a = std::min<int64_t>(rows, per * w), b = std::min<int64_t>(rows, per * (w + 1));
            if (a < b) sgb_unstage_states(src + a, b - a, state_out, out_type, r0 + a);
        });
    }
    return 0;
}

// ---------------------------------------------------------------- viterbi / logspace_add / utilities
int sgb_viterbi(const double* emission, int64_t nrow_e, const double* transition, const int32_t* observations,
                int64_t T, int32_t S, int32_t init_mode, int32_t decode, int32_t* path_out)
{
    if (T < 0 || nrow_e < 0) return fail("negative size");
    if (S < 1 || S > 128) return fail("viterbi: number of states must be in 1..128");
    if (T == 0) return 0;
    if (!emission || !transition || !observations || !path_out) return fail("NULL pointer");
    if (decode != SGB_DECODE_REFERENCE && decode != SGB_DECODE_CHAINED) return fail("decode invalid");
    for (int64_t t = 0; t < T; t++)
        if (observations[t] < 1 || observations[t] > nrow_e) return fail("observations index outside the emission matrix");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    size_t eb = (size_t)nrow_e * S * 8, tb = (size_t)S * S * 8, ob = (size_t)T * 4;
    if (C->scratch_a.reserve(eb + tb + 64) || C->scratch_b.reserve(2 * ob + 64)) return 1;
    if (decode == SGB_DECODE_CHAINED && C->scratch_c.reserve(sgbk_viterbi_bp_bytes(T, S) + 64)) return 1;
    double* d_em = (double*)C->scratch_a.p;
    double* d_tr = d_em + (size_t)nrow_e * S;
    int32_t* d_obs = (int32_t*)C->scratch_b.p;
    int32_t* d_path = d_obs + T;
    SGB_CUDA(cudaMemcpyAsync(d_em, emission, eb, cudaMemcpyHostToDevice, st));
    SGB_CUDA(cudaMemcpyAsync(d_tr, transition, tb, cudaMemcpyHostToDevice, st));
    SGB_CUDA(cudaMemcpyAsync(d_obs, observations, ob, cudaMemcpyHostToDevice, st));
    SGB_CUDA(sgbk_viterbi(d_em, nrow_e, d_tr, d_obs, T, S, init_mode, decode, (uint8_t*)C->scratch_c.p, d_path, st));
    SGB_CUDA(cudaMemcpyAsync(path_out, d_path, ob, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int sgb_logspace_add_vec(const double* logx, const double* logy, int64_t n, double* out)
{
    if (n < 0) return fail("negative size");
    if (n == 0) return 0;
    if (!logx || !logy || !out) return fail("NULL pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    if (C->scratch_a.reserve((size_t)n * 24 + 64)) return 1;
    double* d_a = (double*)C->scratch_a.p;
    double* d_b = d_a + n;
    double* d_o = d_b + n;
    SGB_CUDA(cudaMemcpyAsync(d_a, logx, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    SGB_CUDA(cudaMemcpyAsync(d_b, logy, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    SGB_CUDA(sgbk_logspace_add(d_a, d_b, n, d_o, st));
    SGB_CUDA(cudaMemcpyAsync(out, d_o, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int sgb_logspace_add(double logx, double logy, double* out)
{
    if (!out) return fail("NULL pointer");
    return sgb_logspace_add_vec(&logx, &logy, 1, out);
}

int sgb_math_eval(int32_t fn, const double* in, const double* in2, int64_t n, double* out)
{
    if (n <= 0) return 0;
    if (!in || !out) return fail("NULL pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    if (C->scratch_a.reserve((size_t)n * 24 + 64)) return 1;
    double* d_a = (double*)C->scratch_a.p;
    double* d_b = d_a + n;
    double* d_o = d_b + n;
    SGB_CUDA(cudaMemcpyAsync(d_a, in, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    SGB_CUDA(cudaMemcpyAsync(d_b, in2 ? in2 : in, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    SGB_CUDA(sgbk_math_eval(fn, d_a, d_b, n, d_o, st));
    SGB_CUDA(cudaMemcpyAsync(out, d_o, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int sgb_postprocess(int64_t N, const void* cn, int32_t cn_itemsize, const uint8_t* path, int32_t* A, int32_t* B,
                    int32_t* state_min, uint8_t* loh, uint8_t* phase, uint8_t* state_phase, double* state_baf)
{
    if (N < 0) return fail("negative size");
    if (N == 0) return 0;
    if (cn_itemsize != 1 && cn_itemsize != 4) return fail("cn_itemsize must be 1 (uint8) or 4 (int32)");
    if (!cn || !path || !A || !B || !state_min || !loh || !phase || !state_phase || !state_baf) return fail("NULL pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    // chunked so the scratch stays modest for very long tables
    const int64_t chunk = 16ll << 20;
    if (C->scratch_a.reserve((size_t)std::min(N, chunk) * (cn_itemsize + 1) + 64)) return 1;
    if (C->scratch_b.reserve((size_t)std::min(N, chunk) * 12 + 64)) return 1;
    if (C->scratch_c.reserve((size_t)std::min(N, chunk) * 3 + 64)) return 1;
    if (C->scratch_d.reserve((size_t)std::min(N, chunk) * 8 + 64)) return 1;
    for (int64_t o = 0; o < N; o += chunk) {
        const int64_t m = std::min(chunk, N - o);
        char* d_cn = (char*)C->scratch_a.p;
        uint8_t* d_path = (uint8_t*)d_cn + (size_t)m * cn_itemsize;
        int32_t* d_A = (int32_t*)C->scratch_b.p;
        int32_t* d_B = d_A + m;
        int32_t* d_sm = d_B + m;
        uint8_t* d_loh = (uint8_t*)C->scratch_c.p;
        uint8_t* d_ph = d_loh + m;
        uint8_t* d_sp = d_ph + m;
        double* d_baf = (double*)C->scratch_d.p;
        SGB_CUDA(cudaMemcpyAsync(d_cn, (const char*)cn + (size_t)o * cn_itemsize, (size_t)m * cn_itemsize, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(d_path, path + o, (size_t)m, cudaMemcpyHostToDevice, st));
        SGB_CUDA(sgbk_postprocess(d_cn, cn_itemsize, d_path, m, d_A, d_B, d_sm, d_loh, d_ph, d_sp, d_baf, st));
        SGB_CUDA(cudaMemcpyAsync(A + o, d_A, (size_t)m * 4, cudaMemcpyDeviceToHost, st));
        SGB_CUDA(cudaMemcpyAsync(B + o, d_B, (size_t)m * 4, cudaMemcpyDeviceToHost, st));
        SGB_CUDA(cudaMemcpyAsync(state_min + o, d_sm, (size_t)m * 4, cudaMemcpyDeviceToHost, st));
        SGB_CUDA(cudaMemcpyAsync(loh + o, d_loh, (size_t)m, cudaMemcpyDeviceToHost, st));
        SGB_CUDA(cudaMemcpyAsync(phase + o, d_ph, (size_t)m, cudaMemcpyDeviceToHost, st));
        SGB_CUDA(cudaMemcpyAsync(state_phase + o, d_sp, (size_t)m, cudaMemcpyDeviceToHost, st));
        SGB_CUDA(cudaMemcpyAsync(state_baf + o, d_baf, (size_t)m * 8, cudaMemcpyDeviceToHost, st));
        SGB_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}

int sgb_pass_stats(int64_t N, const int32_t* n, const int32_t* x, const void* cn, int32_t cn_itemsize,
                   const uint8_t* path, int32_t n_states, int64_t* sums_out, int64_t* ab_hist_out)
{
    if (N < 0) return fail("negative size");
    if (cn_itemsize != 1 && cn_itemsize != 4) return fail("cn_itemsize must be 1 (uint8) or 4 (int32)");
    if (n_states < 1 || n_states > SGB_MAX_STATES) return fail("n_states (maxCN+1) must be in 1..32");
    if (!sums_out || !ab_hist_out) return fail("NULL pointer");
    const int S = n_states;
    const size_t nout = 3 + (size_t)S * S;
    for (int i = 0; i < 3; i++) sums_out[i] = 0;
    for (size_t i = 0; i < (size_t)S * S; i++) ab_hist_out[i] = 0;
    if (N == 0) return 0;
    if (!n || !x || !cn || !path) return fail("NULL pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    const int64_t chunk = 16ll << 20;
    const int64_t m0 = std::min(N, chunk);
    if (C->scratch_a.reserve((size_t)m0 * 8 + 64) || C->scratch_b.reserve((size_t)m0 * (cn_itemsize + 1) + 64) ||
        C->scratch_c.reserve(nout * 8 + 64))
        return 1;
    unsigned long long* d_out = (unsigned long long*)C->scratch_c.p;
    SGB_CUDA(cudaMemsetAsync(d_out, 0, nout * 8, st));
    for (int64_t o = 0; o < N; o += chunk) {
        const int64_t m = std::min(chunk, N - o);
        int32_t* d_n = (int32_t*)C->scratch_a.p;
        int32_t* d_x = d_n + m;
        char* d_cn = (char*)C->scratch_b.p;
        uint8_t* d_path = (uint8_t*)d_cn + (size_t)m * cn_itemsize;
        SGB_CUDA(cudaMemcpyAsync(d_n, n + o, (size_t)m * 4, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(d_x, x + o, (size_t)m * 4, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(d_cn, (const char*)cn + (size_t)o * cn_itemsize, (size_t)m * cn_itemsize, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(d_path, path + o, (size_t)m, cudaMemcpyHostToDevice, st));
        SGB_CUDA(sgbk_pass_stats(d_n, d_x, d_cn, cn_itemsize, d_path, m, S, d_out, st));
        SGB_CUDA(cudaStreamSynchronize(st));
    }
    std::vector<unsigned long long> h(nout);
    SGB_CUDA(cudaMemcpy(h.data(), d_out, nout * 8, cudaMemcpyDeviceToHost));
    for (int i = 0; i < 3; i++) sums_out[i] = (int64_t)h[(size_t)i];
    for (size_t i = 0; i < (size_t)S * S; i++) ab_hist_out[i] = (int64_t)h[3 + i];
    return 0;
}

// ---------------------------------------------------------------- table passes
static double g_table_last_ms = 0.0;      // device time of the kernels of the last sgb_table_order / sgb_combine_bafcn call

int sgb_table_last_ms(double* ms_out)
{
    if (!ms_out) return fail("NULL pointer");
    *ms_out = g_table_last_ms;
    return 0;
}

static int table_run(int do_combine, int64_t N, const int32_t* cell, const int32_t* chr, const int32_t* start, const int32_t* state,
                     const int32_t* a, const int32_t* b, const int32_t* tot, int32_t n_cells, int32_t n_chr, double filtern,
                     int32_t minbins, int32_t minbinschr, int32_t chrX, int64_t* nrow_out, int32_t* o_first, int32_t* o_cell, int32_t* o_chr,
                     int32_t* o_start, int32_t* o_state, int32_t* o_a, int32_t* o_b, int32_t* o_n, int64_t* nseg_out, int64_t* seg_off_out)
{
    if (N < 0) return fail("negative size");
    if (N >= (1ll << 31)) return fail("tables of 2^31 rows or more are not supported");
    if (n_cells < 1 || n_chr < 1 || n_chr > 256) return fail("n_cells must be >= 1 and n_chr in 1..256");
    if (!nrow_out || !nseg_out || !seg_off_out) return fail("NULL pointer");
    *nrow_out = 0; *nseg_out = 0; seg_off_out[0] = 0;
    if (N == 0) return 0;
    if (!cell || !chr || !start || !o_first) return fail("NULL pointer");
    if (do_combine && (!state || !a || !b || !tot || !o_cell || !o_chr || !o_start || !o_state || !o_a || !o_b || !o_n)) return fail("NULL pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    const size_t n4 = ((size_t)N * 4 + 255) & ~(size_t)255;
    const int ncol_in = do_combine ? 7 : 3, ncol_out = do_combine ? 8 : 1;
    const size_t seg_bytes = ((size_t)N + 1) * 8;
    if (C->scratch_a.reserve(n4 * (ncol_in + ncol_out) + seg_bytes + 512)) return 1;
    if (C->scratch_b.reserve(sgbk_table_scratch_bytes(N, n_cells, n_chr))) return 1;
    char* base = (char*)C->scratch_a.p;
    const int32_t* hin[7] = {cell, chr, start, state, a, b, tot};
    int32_t* din[7] = {nullptr};
    for (int k = 0; k < ncol_in; k++) {
        din[k] = (int32_t*)(base + n4 * k);
        SGB_CUDA(cudaMemcpyAsync(din[k], hin[k], (size_t)N * 4, cudaMemcpyHostToDevice, st));
    }
    int32_t* dout[8];
    for (int k = 0; k < ncol_out; k++) dout[k] = (int32_t*)(base + n4 * (ncol_in + k));
    long long* dseg = (long long*)(base + n4 * (ncol_in + ncol_out));
    SGB_CUDA(cudaMemsetAsync(C->d_err, 0, sizeof(int), st));
    long long counts[2] = {0, 0};
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    SGB_CUDA(cudaEventCreate(&t0));
    SGB_CUDA(cudaEventCreate(&t1));
    SGB_CUDA(cudaEventRecord(t0, st));
    SGB_CUDA(sgbk_table(N, din[0], din[1], din[2], din[3], din[4], din[5], din[6], n_cells, n_chr, do_combine, filtern, minbins,
                        minbinschr, chrX, C->scratch_b.p, dout[0], do_combine ? dout[1] : nullptr, do_combine ? dout[2] : nullptr,
                        do_combine ? dout[3] : nullptr, do_combine ? dout[4] : nullptr, do_combine ? dout[5] : nullptr,
                        do_combine ? dout[6] : nullptr, do_combine ? dout[7] : nullptr, dseg, counts, C->d_err, st));
    SGB_CUDA(cudaEventRecord(t1, st));
    SGB_CUDA(cudaEventSynchronize(t1));
    {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t0, t1) == cudaSuccess) g_table_last_ms = (double)ms;
        cudaEventDestroy(t0); cudaEventDestroy(t1);
    }
    int bits = 0;
    SGB_CUDA(cudaMemcpy(&bits, C->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (bits) return fail(err_bits_text(bits));
    int32_t* hout[8] = {o_first, o_cell, o_chr, o_start, o_state, o_a, o_b, o_n};
    for (int k = 0; k < ncol_out; k++)
        if (counts[0] > 0) SGB_CUDA(cudaMemcpyAsync(hout[k], dout[k], (size_t)counts[0] * 4, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaMemcpyAsync(seg_off_out, dseg, ((size_t)counts[1] + 1) * 8, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaStreamSynchronize(st));
    if (counts[1] == 0) seg_off_out[0] = 0;
    *nrow_out = counts[0]; *nseg_out = counts[1];
    return 0;
}

int sgb_table_order(int64_t N, const int32_t* cell, const int32_t* chr, const int32_t* start, int32_t n_cells, int32_t n_chr,
                    int32_t* perm_out, int64_t* nseg_out, int64_t* seg_off_out)
{
    int64_t nrow = 0;
    return table_run(0, N, cell, chr, start, nullptr, nullptr, nullptr, nullptr, n_cells, n_chr, 0.0, 0, 0, -1, &nrow, perm_out, nullptr,
                     nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nseg_out, seg_off_out);
}

int sgb_combine_bafcn(int64_t N, const int32_t* cell, const int32_t* chr, const int32_t* start, const int32_t* state,
                      const int32_t* alleleA, const int32_t* alleleB, const int32_t* totalcounts, int32_t n_cells, int32_t n_chr,
                      double filtern, int32_t minbins, int32_t minbinschr, int32_t chrX_code, int64_t* nrow_out, int32_t* first_row_out,
                      int32_t* cell_out, int32_t* chr_out, int32_t* start_out, int32_t* state_out, int32_t* alleleA_out,
                      int32_t* alleleB_out, int32_t* totalcounts_out, int64_t* nseg_out, int64_t* seg_off_out)
{
    if (filtern != filtern) return fail("NaN parameter");
    return table_run(1, N, cell, chr, start, state, alleleA, alleleB, totalcounts, n_cells, n_chr, filtern, minbins, minbinschr, chrX_code,
                     nrow_out, first_row_out, cell_out, chr_out, start_out, state_out, alleleA_out, alleleB_out, totalcounts_out, nseg_out,
                     seg_off_out);
}

int sgb_pass_stats2(int64_t N, const int32_t* n, const int32_t* x, const void* cn, int32_t cn_itemsize, const uint8_t* path,
                    const uint8_t* row_keep, int32_t n_states, int32_t selA, int32_t selB, double* out9, int64_t* ab_hist_out)
{
    if (N < 0) return fail("negative size");
    if (cn_itemsize != 1 && cn_itemsize != 4) return fail("cn_itemsize must be 1 (uint8) or 4 (int32)");
    if (n_states < 1 || n_states > SGB_MAX_STATES) return fail("n_states (maxCN+1) must be in 1..32");
    if (!out9 || !ab_hist_out) return fail("NULL pointer");
    const int S = n_states;
    for (int i = 0; i < 9; i++) out9[i] = 0.0;
    for (int i = 0; i < S * S; i++) ab_hist_out[i] = 0;
    if (N == 0) return 0;
    if (!n || !x || !cn || !path) return fail("NULL pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    const size_t n4 = ((size_t)N * 4 + 255) & ~(size_t)255, n1 = ((size_t)N + 255) & ~(size_t)255;
    const int nb = sgbk_pass_stats2_blocks(N);
    if (C->scratch_a.reserve(3 * n4 + 2 * n1 + 512) || C->scratch_c.reserve(((size_t)nb * 9 + 16) * 8 + (size_t)S * S * 8 + 512)) return 1;
    char* base = (char*)C->scratch_a.p;
    int32_t* d_n = (int32_t*)base; int32_t* d_x = (int32_t*)(base + n4); char* d_cn = base + 2 * n4;
    uint8_t* d_path = (uint8_t*)(base + 3 * n4); uint8_t* d_keep = d_path + n1;
    double* d_part = (double*)C->scratch_c.p; double* d_out = d_part + (size_t)nb * 9;
    unsigned long long* d_hist = (unsigned long long*)(d_out + 16);
    SGB_CUDA(cudaMemcpyAsync(d_n, n, (size_t)N * 4, cudaMemcpyHostToDevice, st));
    SGB_CUDA(cudaMemcpyAsync(d_x, x, (size_t)N * 4, cudaMemcpyHostToDevice, st));
    SGB_CUDA(cudaMemcpyAsync(d_cn, cn, (size_t)N * cn_itemsize, cudaMemcpyHostToDevice, st));
    SGB_CUDA(cudaMemcpyAsync(d_path, path, (size_t)N, cudaMemcpyHostToDevice, st));
    if (row_keep) SGB_CUDA(cudaMemcpyAsync(d_keep, row_keep, (size_t)N, cudaMemcpyHostToDevice, st));
    SGB_CUDA(sgbk_pass_stats2(d_n, d_x, d_cn, cn_itemsize, d_path, row_keep ? d_keep : nullptr, N, S, selA, selB, d_part, d_out, d_hist, st));
    std::vector<unsigned long long> h((size_t)S * S);
    SGB_CUDA(cudaMemcpyAsync(out9, d_out, 9 * 8, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaMemcpyAsync(h.data(), d_hist, (size_t)S * S * 8, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaStreamSynchronize(st));
    for (int i = 0; i < S * S; i++) ab_hist_out[i] = (int64_t)h[(size_t)i];
    return 0;
}

int sgb_qc_per_cell(int64_t N, const int32_t* n, const int32_t* x, const void* cn, int32_t cn_itemsize, const uint8_t* path,
                    const int32_t* cell_of_row, int32_t n_cells, double* avg_out)
{
    if (N < 0 || n_cells < 0) return fail("negative size");
    if (N >= (1ll << 31)) return fail("tables of 2^31 rows or more are not supported");
    if (cn_itemsize != 1 && cn_itemsize != 4) return fail("cn_itemsize must be 1 (uint8) or 4 (int32)");
    if (n_cells == 0) return 0;
    if (!avg_out) return fail("NULL pointer");
    if (N > 0 && (!n || !x || !cn || !path || !cell_of_row)) return fail("NULL pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    const size_t n4 = ((size_t)N * 4 + 255) & ~(size_t)255, n1 = ((size_t)N + 255) & ~(size_t)255;
    if (C->scratch_a.reserve(4 * n4 + n1 + (size_t)n_cells * 8 + 1024) || C->scratch_b.reserve(sgbk_qc_scratch_bytes(N))) return 1;
    char* base = (char*)C->scratch_a.p;
    int32_t* d_n = (int32_t*)base; int32_t* d_x = (int32_t*)(base + n4); char* d_cn = base + 2 * n4;
    int32_t* d_cell = (int32_t*)(base + 3 * n4); uint8_t* d_path = (uint8_t*)(base + 4 * n4);
    double* d_avg = (double*)(base + 4 * n4 + n1);
    if (N > 0) {
        SGB_CUDA(cudaMemcpyAsync(d_n, n, (size_t)N * 4, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(d_x, x, (size_t)N * 4, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(d_cn, cn, (size_t)N * cn_itemsize, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(d_cell, cell_of_row, (size_t)N * 4, cudaMemcpyHostToDevice, st));
        SGB_CUDA(cudaMemcpyAsync(d_path, path, (size_t)N, cudaMemcpyHostToDevice, st));
    }
    SGB_CUDA(cudaMemsetAsync(C->d_err, 0, sizeof(int), st));
    SGB_CUDA(sgbk_qc_per_cell(d_n, d_x, d_cn, cn_itemsize, d_path, d_cell, N, n_cells, C->scratch_b.p, d_avg, C->d_err, st));
    SGB_CUDA(cudaMemcpyAsync(avg_out, d_avg, (size_t)n_cells * 8, cudaMemcpyDeviceToHost, st));
    int bits = 0;
    SGB_CUDA(cudaMemcpy(&bits, C->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    SGB_CUDA(cudaStreamSynchronize(st));
    if (bits) return fail(err_bits_text(bits));
    return 0;
}

int sgb_fix_assignments(int64_t N, const int32_t* state, const int32_t* A, const int32_t* B, const int32_t* alleleA,
                        const int32_t* alleleB, int64_t nseg, const int64_t* seg_off, double loherror, int32_t likelihood, double rho,
                        int32_t* A_out, int32_t* B_out)
{
    if (N < 0 || nseg < 0) return fail("negative size");
    if (likelihood != SGB_LIK_BINOMIAL && likelihood != SGB_LIK_BETABINOMIAL) return fail("Likelihood model must be one of binomial or betabinomial");
    if (loherror != loherror || rho != rho) return fail("NaN parameter");
    if (N == 0) return 0;
    if (!state || !A || !B || !alleleA || !alleleB || !seg_off || !A_out || !B_out) return fail("NULL pointer");
    if (nseg < 1 || seg_off[0] != 0 || seg_off[nseg] != N) return fail("seg_off must cover rows 0..N");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    cudaStream_t st = C->main;
    const size_t n4 = ((size_t)N * 4 + 255) & ~(size_t)255, n1 = ((size_t)N + 255) & ~(size_t)255;
    if (C->scratch_a.reserve(9 * n4 + n1 + 512) || C->scratch_b.reserve(sgbk_fix_scratch_bytes(N))) return 1;
    std::vector<uint8_t> head((size_t)N, 0);
    for (int64_t s = 0; s < nseg; s++) {
        if (seg_off[s + 1] < seg_off[s]) return fail("seg_off must be non-decreasing");
        if (seg_off[s] < N && seg_off[s + 1] > seg_off[s]) head[(size_t)seg_off[s]] = 1;
    }
    char* base = (char*)C->scratch_a.p;
    int32_t* d[9];
    for (int k = 0; k < 9; k++) d[k] = (int32_t*)(base + n4 * k);
    uint8_t* d_head = (uint8_t*)(base + 9 * n4);
    const int32_t* hin[5] = {state, A, B, alleleA, alleleB};
    for (int k = 0; k < 5; k++) SGB_CUDA(cudaMemcpyAsync(d[k], hin[k], (size_t)N * 4, cudaMemcpyHostToDevice, st));
    SGB_CUDA(cudaMemcpyAsync(d_head, head.data(), (size_t)N, cudaMemcpyHostToDevice, st));
    SGB_CUDA(sgbk_fix_assignments(d[0], d[1], d[2], d[3], d[4], d_head, N, loherror, likelihood == SGB_LIK_BETABINOMIAL && rho > 0.0, rho,
                                  d[5], d[6], d[7], d[8], C->scratch_b.p, st));
    SGB_CUDA(cudaMemcpyAsync(A_out, d[7], (size_t)N * 4, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaMemcpyAsync(B_out, d[8], (size_t)N * 4, cudaMemcpyDeviceToHost, st));
    SGB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int sgb_fp64_peak(int32_t iters, double* tflops_out)
{
    if (!tflops_out) return fail("NULL pointer");
    Ctx* C;
    if (get_ctx(&C)) return 1;
    double tf = 0.0;
    SGB_CUDA(sgbk_fp64_peak(iters, C->sm_count, C->main, &tf));
    *tflops_out = tf;
    return 0;
}

}  // extern "C"
