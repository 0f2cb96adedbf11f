// Multi-GPU entry points of the C ABI (include/rtgpu.h): the scene is replicated, the samples of every pixel are
// partitioned over the ranks, and ONE ncclReduce(sum, float32, root 0) per frame over NVLink 5 / NVSwitch combines
// the per-GPU accumulation buffers — the analogue of the reference's mpsc gather of finished pixels
// (src/main.rs:72,101,114).  Paths are independent (src/main.rs:84-106), so no other exchange exists on this path
// and kernels on different GPUs never wait on each other.
//
// Two ways to form the communicator, behind the same RtComm:
//   rt_comm_create      one process per GPU (torchrun, MPI, ...): every process passes the 128-byte id that rank 0
//                       obtained from rt_comm_unique_id and ferried to the others by any means it has
//   rt_comm_create_all  ONE process that owns n devices (ncclCommInitAll): what a Rust main() would call
#include <nccl.h>

#include <chrono>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "rt_internal.hpp"

using namespace rtb;

static_assert(sizeof(ncclUniqueId) == RT_COMM_UNIQUE_ID_BYTES, "RT_COMM_UNIQUE_ID_BYTES must match ncclUniqueId");

// NCCL is bound at RUN time, when the first multi-rank communicator is made (dlopen by SONAME): librtgpu.so itself
// carries no NCCL dependency, so a single-GPU host needs no NCCL installed, and a process that also hosts another
// NCCL user (PyTorch bundles its own libnccl.so.2) ends up with ONE copy — whichever was loaded first.
#include <dlfcn.h>
namespace {
struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;  // why binding failed (empty: bound)
    bool bound = false;
};
NcclApi& nccl() {
    static NcclApi api = [] {
        NcclApi a;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) {
            a.error = std::string("cannot load libnccl.so.2: ") + dlerror();
            return a;
        }
        auto sym = [&](const char* name) -> void* {
            void* p = dlsym(h, name);
            if (!p && a.error.empty()) a.error = std::string("libnccl lacks ") + name;
            return p;
        };
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
        a.CommInitAll = reinterpret_cast<decltype(a.CommInitAll)>(sym("ncclCommInitAll"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
        a.Reduce = reinterpret_cast<decltype(a.Reduce)>(sym("ncclReduce"));
        a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
        a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
        a.bound = a.error.empty();
        return a;
    }();
    return api;
}
}  // namespace
#define NCCL_BOUND()                                                          \
    do {                                                                      \
        if (!nccl().bound) return rt_fail(RT_ERR_NCCL, nccl().error);         \
    } while (0)

#define NCCL_TRY(expr)                                                                                      \
    do {                                                                                                    \
        ncclResult_t r__ = (expr);                                                                          \
        if (r__ != ncclSuccess)                                                                             \
            return rt_fail(RT_ERR_NCCL, std::string(#expr) + ": " + nccl().GetErrorString(r__));           \
    } while (0)

struct RtComm {
    int n_ranks = 0;
    int first_rank = 0;              // global rank of local device 0; local device i is rank first_rank + i
    std::vector<int> devices;        // local CUDA devices
    std::vector<ncclComm_t> comms;   // one communicator per local device
    ~RtComm() {
        for (size_t i = 0; i < comms.size(); i++) {
            if (!comms[i] || !nccl().bound) continue;
            cudaSetDevice(devices[i]);
            nccl().CommDestroy(comms[i]);
        }
    }
};


extern "C" {

int rt_comm_unique_id(uint8_t out_id[RT_COMM_UNIQUE_ID_BYTES]) {
    if (!out_id) return rt_fail(RT_ERR_ARG, "NULL argument");
    NCCL_BOUND();
    ncclUniqueId id;
    NCCL_TRY(nccl().GetUniqueId(&id));
    std::memcpy(out_id, &id, sizeof(id));
    return RT_OK;
}

int rt_comm_create(const uint8_t id_bytes[RT_COMM_UNIQUE_ID_BYTES], int n_ranks, int rank, int device, RtComm** out) {
    if (!out) return rt_fail(RT_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (!id_bytes || n_ranks < 1 || rank < 0 || rank >= n_ranks) return rt_fail(RT_ERR_ARG, "bad communicator arguments");
    int n_dev = 0;
    CU_TRY(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) return rt_fail(RT_ERR_CUDA, "no such CUDA device");
    CU_TRY(cudaSetDevice(device));
    std::unique_ptr<RtComm> c(new RtComm());
    c->n_ranks = n_ranks;
    c->first_rank = rank;
    c->devices = {device};
    c->comms = {nullptr};
    ncclUniqueId id;
    std::memcpy(&id, id_bytes, sizeof(id));
    if (n_ranks > 1) {
        NCCL_BOUND();
        NCCL_TRY(nccl().CommInitRank(&c->comms[0], n_ranks, id, rank));
    }  // (one rank: no collective is ever issued)
    *out = c.release();
    return RT_OK;
}

int rt_comm_create_all(const int* devices, int n_devices, RtComm** out) {
    if (!out) return rt_fail(RT_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (!devices || n_devices < 1) return rt_fail(RT_ERR_ARG, "bad communicator arguments");
    int n_dev = 0;
    CU_TRY(cudaGetDeviceCount(&n_dev));
    for (int i = 0; i < n_devices; i++) {
        if (devices[i] < 0 || devices[i] >= n_dev) return rt_fail(RT_ERR_CUDA, "no such CUDA device");
        for (int j = 0; j < i; j++)
            if (devices[j] == devices[i]) return rt_fail(RT_ERR_ARG, "device listed twice");
    }
    std::unique_ptr<RtComm> c(new RtComm());
    c->n_ranks = n_devices;
    c->first_rank = 0;
    c->devices.assign(devices, devices + n_devices);
    c->comms.assign((size_t)n_devices, nullptr);
    if (n_devices > 1) {
        NCCL_BOUND();
        NCCL_TRY(nccl().CommInitAll(c->comms.data(), n_devices, devices));
    }  // (one rank: no collective is ever issued)
    *out = c.release();
    return RT_OK;
}

int rt_comm_info(const RtComm* comm, int* n_ranks, int* first_rank, int* n_local) {
    if (!comm) return rt_fail(RT_ERR_ARG, "NULL argument");
    if (n_ranks) *n_ranks = comm->n_ranks;
    if (first_rank) *first_rank = comm->first_rank;
    if (n_local) *n_local = (int)comm->devices.size();
    return RT_OK;
}

void rt_comm_destroy(RtComm* comm) { delete comm; }

int rt_render_reduce_device(RtScene* const* scenes, RtComm* comm, const RtRenderParams* frame, float* const* accum, void* const* streams) {
    if (!scenes || !comm || !frame || !accum) return rt_fail(RT_ERR_ARG, "NULL argument");
    const int n_local = (int)comm->devices.size();
    const size_t nf = (size_t)frame->width * frame->height * 3;
    if (nf == 0) return rt_fail(RT_ERR_ARG, "empty image");
    for (int i = 0; i < n_local; i++) {
        if (!scenes[i] || !accum[i]) return rt_fail(RT_ERR_ARG, "NULL scene / buffer");
        if (scenes[i]->device != comm->devices[i]) return rt_fail(RT_ERR_ARG, "scene i must live on the communicator's device i");
    }
    // (1) every local rank renders its share of the frame's sample range into its own zeroed sum buffer
    for (int i = 0; i < n_local; i++) {
        RtScene* s = scenes[i];
        cudaStream_t st = (streams && streams[i]) ? (cudaStream_t)streams[i] : s->stream;
        RtRenderParams rp = *frame;
        int rc = rt_partition_samples(frame->sample_begin, frame->sample_end, comm->n_ranks, comm->first_rank + i, &rp.sample_begin, &rp.sample_end);
        if (rc != RT_OK) return rc;
        CU_TRY(cudaSetDevice(s->device));
        CU_TRY(cudaMemsetAsync(accum[i], 0, nf * sizeof(float), st));
        rc = rt_render_device(s, &rp, accum[i], st, nullptr);
        if (rc != RT_OK) return rc;
    }
    // (2) ONE sum-reduce to rank 0, in place, stream-ordered behind the render kernels
    if (comm->n_ranks > 1) {
        NCCL_TRY(nccl().GroupStart());
        for (int i = 0; i < n_local; i++) {
            cudaStream_t st = (streams && streams[i]) ? (cudaStream_t)streams[i] : scenes[i]->stream;
            ncclResult_t r = nccl().Reduce(accum[i], accum[i], nf, ncclFloat, ncclSum, 0, comm->comms[i], st);
            if (r != ncclSuccess) {
                nccl().GroupEnd();
                return rt_fail(RT_ERR_NCCL, std::string("ncclReduce: ") + nccl().GetErrorString(r));
            }
        }
        NCCL_TRY(nccl().GroupEnd());
    }
    // (3) the root divides by the frame's sample count (c /= samples_per_pixel, src/main.rs:98)
    if (comm->first_rank == 0) {
        RtScene* s = scenes[0];
        cudaStream_t st = (streams && streams[0]) ? (cudaStream_t)streams[0] : s->stream;
        CU_TRY(cudaSetDevice(s->device));
        const float spp = (float)(frame->spp_total ? frame->spp_total : (frame->sample_end - frame->sample_begin));
        CU_TRY(launch_divide(accum[0], nf, spp, st));
        CU_TRY(cudaEventRecord(s->ev_last, st));
    }
    return RT_OK;
}

int rt_render_reduce(RtScene* const* scenes, RtComm* comm, const RtRenderParams* frame, float* out_rgb_mean, RtStats* stats) {
    if (!scenes || !comm || !frame) return rt_fail(RT_ERR_ARG, "NULL argument");
    auto wall0 = std::chrono::steady_clock::now();
    const int n_local = (int)comm->devices.size();
    const size_t nf = (size_t)frame->width * frame->height * 3;
    if (nf == 0) return rt_fail(RT_ERR_ARG, "empty image");
    const bool root = comm->first_rank == 0;
    if (root && !out_rgb_mean) return rt_fail(RT_ERR_ARG, "the process that owns rank 0 must pass out_rgb_mean");
    std::vector<Lease> bufs((size_t)n_local);
    std::vector<float*> accum((size_t)n_local, nullptr);
    for (int i = 0; i < n_local; i++) {
        if (!scenes[i]) return rt_fail(RT_ERR_ARG, "NULL scene");
        CU_TRY(cudaSetDevice(scenes[i]->device));
        CU_TRY(bufs[i].device_mem(scenes[i]->device, nf * sizeof(float)));
        accum[i] = bufs[i].as<float>();
    }
    int rc = rt_render_reduce_device(scenes, comm, frame, accum.data(), nullptr);
    if (rc == RT_OK && root) {
        CU_TRY(cudaSetDevice(scenes[0]->device));
        rc = d2h_frame(scenes[0], scenes[0]->stream, accum[0], out_rgb_mean, nf * sizeof(float));
    }
    RtStats total{};
    for (int i = 0; i < n_local; i++) {  // every local stream is drained before its buffer returns to the cache
        cudaSetDevice(scenes[i]->device);
        cudaError_t e = cudaStreamSynchronize(scenes[i]->stream);
        if (e != cudaSuccess && rc == RT_OK) rc = rt_fail(RT_ERR_CUDA, std::string("render_reduce: ") + cudaGetErrorString(e));
        if (rc == RT_OK && stats) {
            unsigned long long h[2] = {0, 0};
            if (cudaMemcpy(h, scenes[i]->scratch, sizeof(h), cudaMemcpyDeviceToHost) == cudaSuccess) {
                total.rays += h[0];
                total.shade_calls += h[1];
            }
            uint32_t b = 0, e2 = 0;
            rt_partition_samples(frame->sample_begin, frame->sample_end, comm->n_ranks, comm->first_rank + i, &b, &e2);
            total.paths += (uint64_t)frame->width * frame->height * (e2 - b);
            float ms = 0.0f;
            if (e2 > b && cudaEventElapsedTime(&ms, scenes[i]->ev0, scenes[i]->ev1) == cudaSuccess) total.ms_kernel = std::max<double>(total.ms_kernel, ms);
            cudaGetLastError();
            total.launches += (e2 > b ? 1u : 0u);
        }
    }
    if (rc != RT_OK) return rc;
    if (stats) {
        total.launches += root ? 1u : 0u;
        total.mode_used = RT_MODE_CTA_WAVEFRONT;
        total.ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
        *stats = total;
    }
    return RT_OK;
}

}  // extern "C"
