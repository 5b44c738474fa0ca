// Per-device context: one stream, a growable named workspace, pinned staging.  Host-side C++ only
// (no kernels here); compiled into libultragroth_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace ug {

#define UG_CUDA_CHECK(expr)                                                                         \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess) {                                                                    \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + \
                                     __FILE__ + ":" + std::to_string(__LINE__) + " (" #expr ")");    \
        }                                                                                           \
    } while (0)

struct DeviceBuffer {
    void* ptr = nullptr;
    size_t bytes = 0;
};

class DeviceCtx {
public:
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    uint64_t kernel_launches = 0;  // counted by the launch wrappers (bench.py reports it)
    // per-kernel device time, measured with CUDA events on the launching stream when `profile` is on
    bool profile = false;
    struct KernelStat { double ms = 0; uint64_t launches = 0; };
    std::map<std::string, KernelStat> kernel_stats;
    std::map<std::string, uint64_t> counters;  // work counters collected in profile mode (e.g. MSM digit entries)

    explicit DeviceCtx(int dev) : device(dev) {
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0) {
            throw std::runtime_error(std::string("ultragroth_b200: no CUDA device available (") +
                                     (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                                     "); this library has no CPU fallback");
        }
        if (dev < 0 || dev >= count) throw std::runtime_error("ultragroth_b200: invalid device index " + std::to_string(dev));
        UG_CUDA_CHECK(cudaSetDevice(dev));
        cudaDeviceProp prop;
        UG_CUDA_CHECK(cudaGetDeviceProperties(&prop, dev));
        sm_count = prop.multiProcessorCount;
        if (prop.major < 10) {
            throw std::runtime_error("ultragroth_b200: device " + std::string(prop.name) + " is sm_" + std::to_string(prop.major) +
                                     std::to_string(prop.minor) + "; this library is built for sm_100a (B200) only");
        }
        UG_CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        // experiment knob: L2 fetch granularity in bytes (32 / 64 / 128; a device-wide hint).  The bucket accumulation gathers 64-byte
        // points at random and ncu counts ~147 B of DRAM traffic per entry: is that the 128-byte L2 line fill?
        if (const char* g = getenv("UGB_L2_FETCH")) {
            if (cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(g)) != cudaSuccess) cudaGetLastError();
        }
    }
    ~DeviceCtx() {
        cudaSetDevice(device);
        for (auto& kv : ws_) cudaFree(kv.second.ptr);
        for (auto& kv : pinned_) cudaFreeHost(kv.second.ptr);
        for (auto& kv : events_) cudaEventDestroy(kv.second);
        for (auto& p : pending_) { cudaEventDestroy(p.e0); cudaEventDestroy(p.e1); }
        for (auto& e : free_events_) cudaEventDestroy(e);
        if (stream) cudaStreamDestroy(stream);
    }
    DeviceCtx(const DeviceCtx&) = delete;
    DeviceCtx& operator=(const DeviceCtx&) = delete;

    void activate() const { UG_CUDA_CHECK(cudaSetDevice(device)); }

    // named scratch buffer, grown (never shrunk) on demand; contents undefined after growth
    void* workspace(const std::string& name, size_t bytes) {
        DeviceBuffer& b = ws_[name];
        if (b.bytes < bytes) {
            activate();
            if (b.ptr) { UG_CUDA_CHECK(cudaStreamSynchronize(stream)); UG_CUDA_CHECK(cudaFree(b.ptr)); b.ptr = nullptr; b.bytes = 0; }
            size_t want = bytes + bytes / 8;
            cudaError_t e = cudaMalloc(&b.ptr, want);
            if (e != cudaSuccess) { cudaGetLastError(); want = bytes; UG_CUDA_CHECK(cudaMalloc(&b.ptr, want)); }
            b.bytes = want;
        }
        return b.ptr;
    }
    void* pinned(const std::string& name, size_t bytes) {
        DeviceBuffer& b = pinned_[name];
        if (b.bytes < bytes) {
            if (b.ptr) { UG_CUDA_CHECK(cudaStreamSynchronize(stream)); UG_CUDA_CHECK(cudaFreeHost(b.ptr)); b.ptr = nullptr; }
            UG_CUDA_CHECK(cudaMallocHost(&b.ptr, bytes));
            b.bytes = bytes;
        }
        return b.ptr;
    }
    // once-per-context markers (e.g. per-device function attributes already set)
    bool flag(const std::string& name) const { return flags_.count(name) != 0; }
    void set_flag(const std::string& name) { flags_[name] = true; }
    void sync() { activate(); UG_CUDA_CHECK(cudaStreamSynchronize(stream)); }
    // kernel profiling: prof_begin / prof_end bracket one launch; prof_collect folds finished pairs into kernel_stats
    void prof_begin(const char* name) {
        if (!profile) return;
        Pending p;
        p.name = name;
        p.e0 = get_event();
        p.e1 = get_event();
        UG_CUDA_CHECK(cudaEventRecord(p.e0, stream));
        pending_.push_back(p);
    }
    void prof_end() {
        if (!profile || pending_.empty()) return;
        UG_CUDA_CHECK(cudaEventRecord(pending_.back().e1, stream));
    }
    void prof_collect() {
        if (pending_.empty()) return;
        UG_CUDA_CHECK(cudaStreamSynchronize(stream));
        for (auto& p : pending_) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, p.e0, p.e1) == cudaSuccess) {
                KernelStat& st = kernel_stats[p.name];
                st.ms += ms;
                st.launches++;
            } else {
                cudaGetLastError();
            }
            free_events_.push_back(p.e0);
            free_events_.push_back(p.e1);
        }
        pending_.clear();
    }
    // named event recorded on the stream / waited on by the host
    void record(const std::string& name) {
        cudaEvent_t& ev = events_[name];
        if (!ev) UG_CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        UG_CUDA_CHECK(cudaEventRecord(ev, stream));
    }
    void wait(const std::string& name) {
        auto it = events_.find(name);
        if (it == events_.end()) throw std::runtime_error("ultragroth_b200: unknown event " + name);
        UG_CUDA_CHECK(cudaEventSynchronize(it->second));
    }

private:
    struct Pending { const char* name; cudaEvent_t e0, e1; };
    std::vector<Pending> pending_;
    std::vector<cudaEvent_t> free_events_;
    cudaEvent_t get_event() {
        if (!free_events_.empty()) { cudaEvent_t e = free_events_.back(); free_events_.pop_back(); return e; }
        cudaEvent_t e;
        UG_CUDA_CHECK(cudaEventCreate(&e));
        return e;
    }
    std::map<std::string, DeviceBuffer> ws_;
    std::map<std::string, DeviceBuffer> pinned_;
    std::map<std::string, cudaEvent_t> events_;
    std::map<std::string, bool> flags_;
};

#define UG_LAUNCH_CHECK(ctx)                \
    do {                                    \
        (ctx).kernel_launches++;            \
        UG_CUDA_CHECK(cudaGetLastError());  \
    } while (0)

// UG_KERNEL(ctx, "name", kernel<<<...>>>(...)): counted, error checked and (in profile mode) event timed
#define UG_KERNEL(ctx, name, ...)   \
    do {                            \
        (ctx).prof_begin(name);     \
        __VA_ARGS__;                \
        (ctx).prof_end();           \
        UG_LAUNCH_CHECK(ctx);       \
    } while (0)

}  // namespace ug
