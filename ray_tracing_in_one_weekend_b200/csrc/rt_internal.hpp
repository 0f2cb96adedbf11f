// Internals shared by librtgpu's translation units (rtgpu.cu, rt_comm.cu, test_hooks.cu): the scene handle, the
// error channel and the process-wide buffer cache.  Nothing here is part of the C ABI (include/rtgpu.h).
#pragma once
#include <cuda_runtime.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/rtgpu.h"
#include "rt_device.cuh"

namespace rtb {

// Sets the thread-local message rt_last_error() returns and hands `code` back.
int rt_fail(int code, const std::string& msg);

#define CU_TRY(expr)                                                                                        \
    do {                                                                                                    \
        cudaError_t e__ = (expr);                                                                           \
        if (e__ != cudaSuccess)                                                                             \
            return ::rtb::rt_fail(RT_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));       \
    } while (0)

// Process-wide cache of device buffers and pinned host buffers.  A host that creates and destroys scenes per frame
// (bench.py's e2e leg does, on purpose) otherwise pays the driver allocator on every call: cudaMalloc/cudaFree are
// device-synchronising and a first teardown costs hundreds of milliseconds.  Buffers are handed out best-fit and
// taken back on release; rt_release_cached_memory() frees what is idle.  Leaky singleton: never destroyed, so no
// CUDA call runs during static destruction.
class BufCache {
public:
    static BufCache& get();
    cudaError_t take_device(int device, size_t bytes, void** out, size_t* got);
    void give_device(int device, void* p, size_t bytes);
    cudaError_t take_pinned(size_t bytes, void** out, size_t* got);
    void give_pinned(void* p, size_t bytes);
    void trim();

private:
    struct Buf {
        void* p;
        size_t bytes;
        int device;  // -1: pinned host
    };
    std::mutex mu_;
    std::vector<Buf> idle_;
    size_t idle_bytes_dev_ = 0, idle_bytes_pin_ = 0;
    cudaError_t take(int device, size_t bytes, void** out, size_t* got);
    void give(int device, void* p, size_t bytes);
};

// RAII lease of a cached buffer.
struct Lease {
    void* p = nullptr;
    size_t bytes = 0;
    int device = -2;  // -2: empty, -1: pinned host, >= 0: device memory
    Lease() = default;
    Lease(const Lease&) = delete;
    Lease& operator=(const Lease&) = delete;
    ~Lease() { release(); }
    cudaError_t device_mem(int dev, size_t n) {
        release();
        cudaError_t e = BufCache::get().take_device(dev, n, &p, &bytes);
        if (e == cudaSuccess) device = dev;
        return e;
    }
    cudaError_t pinned(size_t n) {
        release();
        cudaError_t e = BufCache::get().take_pinned(n, &p, &bytes);
        if (e == cudaSuccess) device = -1;
        return e;
    }
    void release() {
        if (device == -1) BufCache::get().give_pinned(p, bytes);
        else if (device >= 0) BufCache::get().give_device(device, p, bytes);
        p = nullptr;
        bytes = 0;
        device = -2;
    }
    template <class T>
    T* as() const { return reinterpret_cast<T*>(p); }
};

// Launch plan of the CTA-wavefront kernel for one scene: resolved ONCE (first use), never per frame.
struct CtaPlan {
    void* fn = nullptr;
    int threads = 0;
    int blocks_per_sm = 0;
    size_t smem = 0;
    uint32_t n_paths = 0, refill_idle = 0, refill_camera = 0, unified_todo = 0, stage_perlin = 0;
    bool flat = false;  // the warp-uniform primitive loop (tiny scenes) instead of the BVH traversal
};

}  // namespace rtb

struct RtScene;
namespace rtb {
// rtgpu.cu: buf[i] /= divisor on `st` (k_divide), and the device-to-host copy of a frame (direct DMA into
// page-locked destinations, a pinned staging pipeline into pageable ones; waits for the copy).
cudaError_t launch_divide(float* buf, size_t n, float divisor, cudaStream_t st);
int d2h_frame(RtScene* s, cudaStream_t st, const void* dev, void* host, size_t bytes);
}  // namespace rtb

// Opaque scene handle of the C ABI.  One in-flight call per handle (rtgpu.h "Threading").
struct RtScene {
    int device = 0;
    cudaStream_t stream = nullptr;
    rtb::DeviceScene ds{};
    rtb::Lease arena;                      // every table of the scene in ONE device allocation
    void* gpu_nodes = nullptr;             // nodes left on the device by the LBVH builder (cudaMalloc'ed)
    std::vector<cudaTextureObject_t> texobjs;
    RtBvhInfo bvh{};
    RtCamera camera{};
    // render scratch (inside the arena): [0..15] statistics, [16] path counter of the CTA kernel, [17] megakernel work counter
    unsigned long long* scratch = nullptr;
    // Event pairs around the render kernels of the last KERNEL_RING rt_render_device calls (ev0/ev1 = the current
    // pair): rt_scene_kernel_times reads them back without having forced a synchronisation per frame.
    static constexpr int KERNEL_RING = 32;
    cudaEvent_t ring[2 * KERNEL_RING] = {};
    unsigned long long n_renders = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev_last = nullptr;                       // behind the last piece of work any entry point enqueued
    cudaEvent_t ev_stage[2] = {nullptr, nullptr};        // staging pipeline of d2h_frame
    int sm_count = 0;
    uint32_t features = 0;    // F_* bits present in the scene
    uint32_t class_mask = 0;  // shading classes (CLS_*) of the scene's materials
    uint32_t q_cls = 0, q_order = 0;  // CTA wavefront: shade queue -> class (4 bits per queue), drain order
    uint32_t n_world_prims = 0;
    rtb::CtaPlan cta_plan[2];          // [0] plain, [1] with traversal counters
    int mega_blocks_per_sm[2] = {0, 0};
    // development knobs (RT_DEV_KNOBS builds read them from the environment ONCE, in rt_scene_create)
    int knob_mode = 0, knob_paths = 0, knob_refill = 8, knob_blocks = 64, knob_carveout = -1, knob_full_kernel = 0,
        knob_units_per_warp = 24, knob_flat = -1, knob_unified = -1, knob_refill_camera = 24;

    ~RtScene();
};
