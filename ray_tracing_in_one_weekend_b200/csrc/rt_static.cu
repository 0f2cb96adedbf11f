// rt_closest_hit_static — the batch closest hit over HOST buffers with NO allocation of any kind.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  rt_closest_hit stages
// the rays and the hit records in device scratch memory it leases from the library's buffer cache (cudaMalloc on a
// miss); this variant keeps that scratch in ONE file-scope `static __device__` array of a stated maximum size
// (RT_STATIC_MAX_RAYS rays = 4.5 MiB), which every GPU that loads the library holds until the process exits.
//
// In HOST code the array's name is a host-side stand-in, not a device address: its address is looked up once per
// device, with that device current, through cudaGetSymbolAddress, and that pointer is what the copies and the kernel
// receive.  The entry point calls no cudaMalloc / cudaFree / allocating library routine.
//
// EVERY call on a device shares the array, and it keeps the last call's contents: the one word that must start at
// zero (the hit counter) is cleared by every call; calls that overlap on one GPU (two host threads, two scenes on
// the same device) race on it — serialise them.
#include <mutex>
#include <string>

#include "rt_internal.hpp"

using namespace rtb;

namespace {

constexpr size_t ST_MAX = RT_STATIC_MAX_RAYS;
constexpr size_t ST_HEADER = 256;                          // [0] u64 hit counter (cleared per call)
constexpr size_t ST_OFF_O = ST_HEADER;                     // n x 3 floats
constexpr size_t ST_OFF_D = ST_OFF_O + ST_MAX * 12;        // n x 3 floats
constexpr size_t ST_OFF_T = ST_OFF_D + ST_MAX * 12;        // n floats
constexpr size_t ST_OFF_H = ST_OFF_T + ST_MAX * 4;         // n x RtHit (44 B)
constexpr size_t ST_BYTES = ST_OFF_H + ST_MAX * sizeof(RtHit);
static_assert(ST_OFF_D % 256 == 0 && ST_OFF_T % 256 == 0 && ST_OFF_H % 256 == 0, "sections stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_static_scratch[ST_BYTES];

// Variant of k_closest_hit_batch (kernels.cuh): the same device functions (closest_hit<F_ALL>, fill_hit<F_ALL>), the
// rays and the records in the static scratch, plus a count of the rays that hit something.
__global__ void __launch_bounds__(128) k_closest_hit_static(DeviceScene sc, unsigned char* scratch, int has_times, uint32_t n,
                                                            float t_min, float t_max) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const float* o_xyz = reinterpret_cast<const float*>(scratch + ST_OFF_O);
    const float* d_xyz = reinterpret_cast<const float*>(scratch + ST_OFF_D);
    const float* times = reinterpret_cast<const float*>(scratch + ST_OFF_T);
    RtHit* out = reinterpret_cast<RtHit*>(scratch + ST_OFF_H);
    bool hit = false;
    if (i < n) {
        Ray r;
        r.o = mk3(o_xyz[3 * i], o_xyz[3 * i + 1], o_xyz[3 * i + 2]);
        r.d = mk3(d_xyz[3 * i], d_xyz[3 * i + 1], d_xyz[3 * i + 2]);
        r.tm = has_times ? times[i] : 0.0f;
        PcgRng rng;
        rng.state = 0;
        const ClosestHit h = closest_hit<F_ALL, false>(sc, r, t_min, t_max, rng, /*skip_media=*/true, nullptr);
        RtHit o;
        o.prim = -1;
        o.t = 0.0f;
        o.p[0] = o.p[1] = o.p[2] = 0.0f;
        o.n[0] = o.n[1] = o.n[2] = 0.0f;
        o.uv[0] = o.uv[1] = 0.0f;
        o.front_face = 0u;
        if (h.prim != 0xffffffffu) {
            const HitRec rec = fill_hit<F_ALL>(sc, r, h.prim, h.t, true);
            o.prim = (int32_t)h.flat;
            o.t = rec.t;
            o.p[0] = rec.p.x; o.p[1] = rec.p.y; o.p[2] = rec.p.z;
            o.n[0] = rec.norm.x; o.n[1] = rec.norm.y; o.n[2] = rec.norm.z;
            o.uv[0] = rec.u; o.uv[1] = rec.v;
            o.front_face = rec.front_face ? 1u : 0u;
            hit = true;
        }
        out[i] = o;
    }
    const uint32_t m = __ballot_sync(0xffffffffu, hit);
    if ((threadIdx.x & 31u) == 0u && m != 0u) atomicAdd(reinterpret_cast<unsigned long long*>(scratch), (unsigned long long)__popc(m));
}

// Device address of the array, looked up once per device (with that device current).
constexpr int ST_MAX_DEVICES = 64;
std::mutex g_mu;
unsigned char* g_dev_ptr[ST_MAX_DEVICES] = {};

int scratch_for(int device, unsigned char** out) {
    if (device < 0 || device >= ST_MAX_DEVICES) return rt_fail(RT_ERR_ARG, "rt_closest_hit_static: device index beyond 63");
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_dev_ptr[device]) {
        void* p = nullptr;
        CU_TRY(cudaGetSymbolAddress(&p, g_static_scratch));  // the caller made `device` current
        g_dev_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_dev_ptr[device];
    return RT_OK;
}

}  // namespace

extern "C" int rt_closest_hit_static(RtScene* s, const float* o_xyz, const float* d_xyz, const float* times, size_t n, float t_min,
                                     float t_max, RtHit* out, uint64_t* out_n_hits) {
    if (!s || !o_xyz || !d_xyz || !out) return rt_fail(RT_ERR_ARG, "NULL argument");
    if (n > ST_MAX)
        return rt_fail(RT_ERR_ARG, "rt_closest_hit_static: " + std::to_string(n) + " rays exceed RT_STATIC_MAX_RAYS (" +
                                       std::to_string(ST_MAX) + "): split the batch or call rt_closest_hit");
    if (out_n_hits) *out_n_hits = 0;
    if (n == 0) return RT_OK;
    CU_TRY(cudaSetDevice(s->device));
    unsigned char* scratch = nullptr;
    int rc = scratch_for(s->device, &scratch);
    if (rc != RT_OK) return rc;
    cudaStream_t st = s->stream;
    CU_TRY(cudaMemsetAsync(scratch, 0, ST_HEADER, st));  // the hit counter starts at zero; everything else is overwritten
    CU_TRY(cudaMemcpyAsync(scratch + ST_OFF_O, o_xyz, n * 12, cudaMemcpyHostToDevice, st));
    CU_TRY(cudaMemcpyAsync(scratch + ST_OFF_D, d_xyz, n * 12, cudaMemcpyHostToDevice, st));
    if (times) CU_TRY(cudaMemcpyAsync(scratch + ST_OFF_T, times, n * 4, cudaMemcpyHostToDevice, st));
    k_closest_hit_static<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(s->ds, scratch, times ? 1 : 0, (uint32_t)n, t_min, t_max);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaEventRecord(s->ev_last, st));
    unsigned long long n_hits = 0;
    CU_TRY(cudaMemcpyAsync(out, scratch + ST_OFF_H, n * sizeof(RtHit), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(&n_hits, scratch, sizeof(n_hits), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (out_n_hits) *out_n_hits = n_hits;
    return RT_OK;
}
