// host_stage.h -- host-side staging of the caller's columns for the chunked host->device pipeline (capi.cu).
//
// The R package hands over pageable INTSXP (int32) or REALSXP (double) vectors (R/callHSCN.R:244-280 works on a
// data.table).  A cudaMemcpyAsync from pageable memory is staged by the driver on ONE thread and blocks the caller, and
// 12-24 bytes per cell-bin cross PCIe.  Instead a small pool of worker threads copies every chunk into page-locked
// staging buffers while NARROWING it to what the kernels need -- read counts to uint16 when the chunk's values allow
// (DLP bins hold a few thousand reads at most; otherwise int32), copy-number states to uint8 -- so 5 bytes per cell-bin
// cross the bus, and converts the decoded states back into the caller's element type.  This is data movement only:
// no likelihood or state is computed on the host.
#pragma once
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#define SGB_T_U8 1
#define SGB_T_U16 2
#define SGB_T_I32 4
#define SGB_T_F64 8

static inline int sgb_type_size(int t) { return t == SGB_T_U8 ? 1 : t == SGB_T_U16 ? 2 : t == SGB_T_I32 ? 4 : t == SGB_T_F64 ? 8 : 0; }

// result bits of a staging pass over one range
#define SGB_STG_WIDE 1          // a count does not fit uint16 (or is negative): restage the chunk as int32, the device reports it
#define SGB_STG_NOT_INT 2       // a double column holds NaN / a non-integer / a value beyond int32
#define SGB_STG_BAD_CN 4        // copy number outside 0..255

class SgbPool {
public:
    explicit SgbPool(int n) : n_(n < 1 ? 1 : n)
    {
        for (int i = 1; i < n_; i++) th_.emplace_back([this, i] { loop(i); });
    }
    ~SgbPool()
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
            gen_++;
        }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    int size() const { return n_; }
    // runs job(worker, nworkers) on every worker (the caller is worker 0) and returns when all are done
    void run(const std::function<void(int, int)>& job)
    {
        if (n_ == 1) { job(0, 1); return; }
        {
            std::lock_guard<std::mutex> lk(mu_);
            job_ = &job;
            pending_ = n_ - 1;
            gen_++;
        }
        cv_.notify_all();
        job(0, n_);
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [this] { return pending_ == 0; });
        job_ = nullptr;
    }

private:
    void loop(int id)
    {
        unsigned long long seen = 0;
        for (;;) {
            const std::function<void(int, int)>* job;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                job = job_;
            }
            if (job) (*job)(id, n_);
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }
    int n_;
    std::vector<std::thread> th_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    const std::function<void(int, int)>* job_ = nullptr;
    unsigned long long gen_ = 0;
    int pending_ = 0;
    bool stop_ = false;
};

// ---- conversion loops (plain loops the compiler vectorises; `bad` accumulates SGB_STG_* bits) --------------------
#if defined(__GNUC__) && !defined(__clang__)
#define SGB_STAGE_OPT __attribute__((optimize("O3", "tree-vectorize")))
#else
#define SGB_STAGE_OPT
#endif

SGB_STAGE_OPT static inline int sgb_stage_count_u16(const void* src, int type, int64_t a, int64_t b, uint16_t* dst)
{
    unsigned wide = 0, notint = 0;
    if (type == SGB_T_I32) {
        const int32_t* s = (const int32_t*)src;
        for (int64_t i = a; i < b; i++) { const int32_t v = s[i]; wide |= (uint32_t)v >> 16; dst[i - a] = (uint16_t)v; }
    } else if (type == SGB_T_F64) {
        const double* s = (const double*)src;
        for (int64_t i = a; i < b; i++) {
            const double v = s[i];
            const bool in = v >= 0.0 && v <= 65535.0;               // false for NaN
            const int32_t iv = in ? (int32_t)v : 0;
            wide |= in ? 0u : 1u;
            notint |= (in && (double)iv != v) ? 1u : 0u;
            dst[i - a] = (uint16_t)iv;
        }
    } else if (type == SGB_T_U16) {
        memcpy(dst, (const uint16_t*)src + a, (size_t)(b - a) * 2);
    } else {
        const uint8_t* s = (const uint8_t*)src;
        for (int64_t i = a; i < b; i++) dst[i - a] = s[i];
    }
    return (wide ? SGB_STG_WIDE : 0) | (notint ? SGB_STG_NOT_INT : 0);
}

SGB_STAGE_OPT static inline int sgb_stage_count_i32(const void* src, int type, int64_t a, int64_t b, int32_t* dst)
{
    unsigned notint = 0;
    if (type == SGB_T_I32) {
        memcpy(dst, (const int32_t*)src + a, (size_t)(b - a) * 4);
    } else if (type == SGB_T_F64) {
        const double* s = (const double*)src;
        for (int64_t i = a; i < b; i++) {
            const double v = s[i];
            const bool in = v >= -2147483648.0 && v <= 2147483647.0;
            const int32_t iv = in ? (int32_t)v : 0;
            notint |= (!in || (double)iv != v) ? 1u : 0u;
            dst[i - a] = iv;
        }
    } else if (type == SGB_T_U16) {
        const uint16_t* s = (const uint16_t*)src;
        for (int64_t i = a; i < b; i++) dst[i - a] = s[i];
    } else {
        const uint8_t* s = (const uint8_t*)src;
        for (int64_t i = a; i < b; i++) dst[i - a] = s[i];
    }
    return notint ? SGB_STG_NOT_INT : 0;
}

SGB_STAGE_OPT static inline int sgb_stage_cn_u8(const void* src, int type, int64_t a, int64_t b, uint8_t* dst)
{
    unsigned badcn = 0, notint = 0;
    if (type == SGB_T_I32) {
        const int32_t* s = (const int32_t*)src;
        for (int64_t i = a; i < b; i++) { const int32_t v = s[i]; badcn |= (uint32_t)v >> 8; dst[i - a] = (uint8_t)v; }
    } else if (type == SGB_T_F64) {
        const double* s = (const double*)src;
        for (int64_t i = a; i < b; i++) {
            const double v = s[i];
            const bool in = v >= 0.0 && v <= 255.0;
            const int32_t iv = in ? (int32_t)v : 0;
            badcn |= (v == v && !in) ? 1u : 0u;
            notint |= (v != v || (in && (double)iv != v)) ? 1u : 0u;
            dst[i - a] = (uint8_t)iv;
        }
    } else if (type == SGB_T_U16) {
        const uint16_t* s = (const uint16_t*)src;
        for (int64_t i = a; i < b; i++) { const uint16_t v = s[i]; badcn |= (unsigned)(v >> 8); dst[i - a] = (uint8_t)v; }
    } else {
        memcpy(dst, (const uint8_t*)src + a, (size_t)(b - a));
    }
    return (badcn ? SGB_STG_BAD_CN : 0) | (notint ? SGB_STG_NOT_INT : 0);
}

// decoded states (uint8) -> the caller's element type
SGB_STAGE_OPT static inline void sgb_unstage_states(const uint8_t* src, int64_t n, void* dst, int type, int64_t at)
{
    if (type == SGB_T_U8) memcpy((uint8_t*)dst + at, src, (size_t)n);
    else if (type == SGB_T_F64) { double* d = (double*)dst + at; for (int64_t i = 0; i < n; i++) d[i] = (double)src[i]; }
    else if (type == SGB_T_I32) { int32_t* d = (int32_t*)dst + at; for (int64_t i = 0; i < n; i++) d[i] = (int32_t)src[i]; }
    else { uint16_t* d = (uint16_t*)dst + at; for (int64_t i = 0; i < n; i++) d[i] = (uint16_t)src[i]; }
}
