// nn_bvh_b200/csrc/engine.cu -- the C ABI of libnnbvh.so (include/nnbvh.h).
//
// Owns device memory, streams and launches.  There is no CPU path in here: if
// CUDA is unavailable every compute entry point returns NNBVH_ERR_CUDA.

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <vector>

#include "../../include/nnbvh.h"
#include "internal.h"
#include "kernels.cuh"
#include "alpha.cuh"
#include "binning.cuh"
#include "hlbvh.cuh"
#include "traverse.cuh"

namespace nnbvh {

static thread_local char g_err[512] = "";

int SetError(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return SetError(NNBVH_ERR_CUDA, "%s failed: %s (%s:%d)", #call,               \
                            cudaGetErrorString(e_), __FILE__, __LINE__);                  \
    } while (0)

template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t cap = 0;  // elements
    int Reserve(size_t n) {
        if (n <= cap) return NNBVH_OK;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        CU(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
        cap = n;
        return NNBVH_OK;
    }
    void Free() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// A cudaMalloc'ed temporary that is freed on every return path.
template <typename T>
struct DevTmp {
    T *p = nullptr;
    DevTmp() = default;
    DevTmp(const DevTmp &) = delete;
    DevTmp &operator=(const DevTmp &) = delete;
    ~DevTmp() {
        if (p) cudaFree(p);
    }
    cudaError_t Alloc(size_t n) { return cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)); }
};

// Known-answer hooks: host array -> device temporary and back, ordered on the engine's stream (a blocking
// cudaMemcpy on the legacy stream is not ordered with a cudaStreamNonBlocking stream).
template <typename T>
static cudaError_t StageIn(DevTmp<T> &d, const void *host, size_t count, cudaStream_t st) {
    cudaError_t ce = d.Alloc(count);
    if (ce != cudaSuccess) return ce;
    return cudaMemcpyAsync(d.p, host, count * sizeof(T), cudaMemcpyHostToDevice, st);
}

// Per-launch device scratch: the persistent kernel's ray cursor and the binning buffers.
constexpr int kCursorWords = 8;
struct Scratch {
    // cursor[0], cursor[1]: ray cursors of the fast and the exact launch; cursor[2] (low
    // 32 bits): number of deferred rays; cursor[3], cursor[4]: binning's distinct-keys / warps
    // sampled; cursor[5] (low 32 bits): binning's "sorted" flag.  Zeroed together before each batch.
    unsigned long long *cursor = nullptr;
    DevBuf<uint32_t> keys, keys2, perm, perm2, hist, tileSums, deferred;
    // Binning mode 2: the device decides per batch, but only the host can skip LAUNCHES.  The flag
    // of every batch is copied here (pinned, asynchronously, never waited for); after two batches
    // in a row that arrived coherent the sort kernels are no longer enqueued, until a batch
    // decides otherwise.  Affects launch overhead only, never results.
    uint32_t *hFlag = nullptr;
    int coherentRun = 0;
    void Free() {
        if (cursor) cudaFree(cursor);
        cursor = nullptr;
        if (hFlag) cudaFreeHost(hFlag);
        hFlag = nullptr;
        keys.Free(); keys2.Free(); perm.Free(); perm2.Free(); hist.Free(); tileSums.Free();
        deferred.Free();
    }
};

// One in-flight chunk of the host-pointer pipeline: H2D rays -> kernel -> D2H results.
struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    cudaEvent_t k0 = nullptr, k1 = nullptr;  // around the chunk's kernels (pooled: created once per slot)
    bool timed = false;                      // k0/k1 hold a chunk whose time has not been added yet
    DevBuf<nnbvh_ray> rays;
    DevBuf<float> soa;       // 7 planes for the SoA entry points
    DevBuf<nnbvh_hit> hits;
    DevBuf<uint8_t> occ;
    DevBuf<uint32_t> cnt;    // 2 planes (nodesVisited, triTests)
    DevBuf<uint32_t> inner;  // instanced scenes: primitive inside the hit object
    Scratch scratch;
};

}  // namespace nnbvh

using namespace nnbvh;

struct SceneStore {
    DevBuf<nnbvh_node> dNodes;
    DevBuf<WideNode> dWide;
    DevBuf<float4> dTris;
    DevBuf<SphereDev> dSpheres;
    DevBuf<InstanceDev> dInstances;
    DevBuf<nnbvh_node> dObjNodes;
    DevBuf<float4> dObjTris;
    DevBuf<float> dAlpha;  // one value per ordered primitive (nnbvh_upload_scene_alpha)
    DevBuf<InstanceStep> dInstStep;  // instanced scenes on the StepKernel
    ~SceneStore() {
        dNodes.Free(); dWide.Free(); dTris.Free(); dSpheres.Free();
        dInstances.Free(); dObjNodes.Free(); dObjTris.Free(); dAlpha.Free(); dInstStep.Free();
    }
};

struct nnbvh_engine {
    int device = 0;
    int kernel = NNBVH_KERNEL_AUTO;
    int smemTopNodes = -1;
    int numSMs = 0;
    int maxSmemOptin = 0;
    cudaStream_t ownStream = nullptr, stream = nullptr;
    cudaEvent_t evA = nullptr, evB = nullptr;  // around a whole batch call (device-pointer path)
    cudaEvent_t evC = nullptr, evD = nullptr;  // around its traversal launches only
    bool haveTraversalEvents = false;
    float lastMs = 0.f;
    uint64_t launches = 0;
    // scene
    bool haveScene = false;
    bool boxesRegular = true;  // every node box finite with lo <= hi: the fast slab test is exact
    SceneDev sc{};
    // device copy of the scene: shared by an engine and its clones (nnbvh_clone), freed with the last of them
    std::shared_ptr<SceneStore> store = std::make_shared<SceneStore>();
    bool hasInstances = false;  // instanced scene: BasicKernel with nested traversal runs
    bool hasSpheres = false;  // scene uploaded with spheres: the SPHERES instantiations run
    bool hasAlpha = false;    // scene uploaded with per-primitive alpha: AlphaKernel runs
    bool general = false;     // nnbvh_upload_scene_general with more than one special kind: BasicKernel<.., 1, 1, 1> runs
    bool instStepOK = false;  // instanced scene laid out for StepKernel<..., INST> (outer + 1 + inner stack depth <= 64)
    bool generalStepOK = false;  // general scene that the StepKernel's general instantiation can run
    // device-pointer path scratch
    Scratch scratch;
    // host-pointer pipeline
    static constexpr int kMaxSlots = 8;
    int nSlots = 4;          // chunks in flight (NNBVH_SLOTS); 512K x 4 measured best on PCIe Gen5
    Slot slots[kMaxSlots];
    // tuning knobs (env overridable, see nnbvh_create)
    int wideMinB = 5;        // CTAs/SM the traversal kernel is compiled for
    // StepKernel lane policy; -1 = by scene size (StepKnobs): the optimum moved down when the node step's child
    // selection became branch-free (r02e-g sweeps: bathroom +3.5 / +2.4 / +8.7 % bounce / shadow / primary at
    // quorum 8, refill 12; the latency-bound scenes -- crown, the synthetic soups -- prefer 10 / 16)
    int refillBelow = -1;
    int leafQuorum = -1;
    int stepStats = 0;       // NNBVH_STEP_STATS=1: run the counting instantiation of StepKernel
    unsigned long long *dStats = nullptr;
    int instStep = 1;        // instanced scenes: 1 = StepKernel<..., INST>, 0 = the one-thread-per-ray BasicKernel
    int gridRaysPerLane = 0; // StepKernel: shrink the persistent grid for small batches (StepGrid); 0 = never
    int sphereQuorum = 12;   // StepKernel with sphere leaves: lanes waiting at spheres before their tests run
    int l2PersistMB = 0;     // NNBVH_L2_PERSIST_MB: L2 set-aside for the wide nodes (access-policy window on the launch stream); 0 = off
    size_t l2PersistBytes = 0, l2MaxWindow = 0;
    int stepBlock = 256;     // StepKernel<.., 5>: threads per CTA (a multiple of 32 up to 256; 192 = 7 CTAs / 42 warps per SM)
    int chunkRamp = 1;       // host-pointer pipeline: small chunks at both ends of a call (RunBatch)
    uint64_t chunkRays = 1u << 19;
    uint64_t chunkRaysSoa = 1u << 20;  // SoA planes are 4 bytes per ray each: larger chunks per copy
    // ray binning pre-pass (binning.cuh)
    int exactOnly = 0;       // skip the fast-slab launch (debug / A-B knob)
    int binMode = 2;         // 0 = off, 1 = always, 2 = per batch, decided on the device (>= binMinRays)
    int binThreshold = 8;    // mode 2: sort when the average warp holds more distinct keys than this
    int binCellBits = -1;    // origin grid (1<<bits)^3; -1 = 5 for cache-resident scenes, 7 beyond (r02i: crown +3.5 %, 10 M soup +3.9 %)
    int binDirBits = 3;      // octahedral direction grid (1<<bits)^2
    uint64_t binMinRays = 1u << 21;  // the pre-pass is ~17 launches: not worth it under ~2 Mi rays
    uint32_t binMinNodes = 1u << 18; // nor for scenes the L1s hold anyway (measured: killeroos loses 11-14 %)
};

// nnbvh_intersect_closest_noalloc's set-up (end of this file): resolves the device address of the
// file-scope scratch array for `device`, which must be current.  Called by nnbvh_create.
static int NoAllocSetup(int device);

// refillBelow / leafQuorum of the StepKernel for the uploaded scene (see nnbvh_engine)
static void StepKnobs(const nnbvh_engine *e, uint32_t *refillBelow, uint32_t *leafQuorum) {
    const size_t sceneBytes = sizeof(WideNode) * (size_t)e->sc.nWide + 48 * (size_t)e->sc.nPrims;
    const bool cacheResident = sceneBytes < ((size_t)96 << 20);  // fits the L2 with room for the ray / hit streams
    const int rb = e->refillBelow > 0 ? e->refillBelow : (cacheResident ? 12 : 16);
    const int lq = e->leafQuorum > 0 ? e->leafQuorum : (cacheResident ? 8 : 10);
    *refillBelow = (uint32_t)std::min(32, std::max(1, rb));
    *leafQuorum = (uint32_t)std::min(32, std::max(1, lq));
}

extern "C" const char *nnbvh_last_error(void) { return g_err; }
extern "C" const char *nnbvh_version(void) { return "nnbvh-b200 0.1 (sm_100a)"; }

static int EnvInt(const char *name, int dflt) {
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

extern "C" int nnbvh_create(const nnbvh_options *opts, nnbvh_engine **out) {
    if (!out) return SetError(NNBVH_ERR_INVALID, "nnbvh_create: out is NULL");
    *out = nullptr;
    int nDev = 0;
    cudaError_t ce = cudaGetDeviceCount(&nDev);
    if (ce != cudaSuccess || nDev == 0)
        return SetError(NNBVH_ERR_CUDA, "nnbvh_create: no CUDA device (%s); this library has "
                                        "no CPU fallback",
                        ce == cudaSuccess ? "device count 0" : cudaGetErrorString(ce));
    // owned until handed to the caller: every early return below destroys what was created so far
    struct Destroyer {
        void operator()(nnbvh_engine *p) const { nnbvh_destroy(p); }
    };
    std::unique_ptr<nnbvh_engine, Destroyer> owner(new nnbvh_engine);
    nnbvh_engine *e = owner.get();
    int dev = opts ? opts->device : -1;
    if (dev < 0) CU(cudaGetDevice(&dev));
    if (dev >= nDev) return SetError(NNBVH_ERR_INVALID, "nnbvh_create: device %d of %d", dev, nDev);
    e->device = dev;
    CU(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10)
        return SetError(NNBVH_ERR_CUDA, "nnbvh_create: device %d is sm_%d%d; this library is "
                                        "built for sm_100a only", dev, prop.major, prop.minor);
    e->numSMs = prop.multiProcessorCount;
    e->maxSmemOptin = (int)prop.sharedMemPerBlockOptin;
    e->kernel = opts ? opts->kernel : NNBVH_KERNEL_AUTO;
    e->smemTopNodes = opts ? opts->smem_top_nodes : -1;
    e->kernel = EnvInt("NNBVH_KERNEL", e->kernel);
    if (e->kernel < NNBVH_KERNEL_AUTO || e->kernel > NNBVH_KERNEL_STEP)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_create: unknown kernel %d", e->kernel);
    e->smemTopNodes = EnvInt("NNBVH_SMEM_NODES", e->smemTopNodes);
    e->wideMinB = EnvInt("NNBVH_WIDE_MINB", e->wideMinB);
    e->refillBelow = EnvInt("NNBVH_REFILL_BELOW", e->refillBelow);
    e->l2PersistMB = std::max(0, EnvInt("NNBVH_L2_PERSIST_MB", e->l2PersistMB));
    if (e->l2PersistMB > 0) {
        int maxPersist = 0, maxWin = 0;
        CU(cudaDeviceGetAttribute(&maxPersist, cudaDevAttrMaxPersistingL2CacheSize, e->device));
        CU(cudaDeviceGetAttribute(&maxWin, cudaDevAttrMaxAccessPolicyWindowSize, e->device));
        e->l2PersistBytes = std::min((size_t)e->l2PersistMB << 20, (size_t)std::max(0, maxPersist));
        e->l2MaxWindow = (size_t)std::max(0, maxWin);
        if (e->l2PersistBytes) CU(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, e->l2PersistBytes));
        if (getenv("NNBVH_VERBOSE")) fprintf(stderr, "nnbvh: L2 persisting max %d B, window max %d B, set-aside %zu B\n", maxPersist, maxWin, e->l2PersistBytes);
    }
    e->stepBlock = std::min(256, std::max(32, EnvInt("NNBVH_STEP_BLOCK", e->stepBlock) & ~31));
    e->leafQuorum = EnvInt("NNBVH_LEAF_QUORUM", e->leafQuorum);
    e->gridRaysPerLane = EnvInt("NNBVH_GRID_RAYS_PER_LANE", e->gridRaysPerLane);
    e->instStep = EnvInt("NNBVH_INST_STEP", e->instStep);
    e->sphereQuorum = EnvInt("NNBVH_SPHERE_QUORUM", e->sphereQuorum);
    e->stepStats = EnvInt("NNBVH_STEP_STATS", 0);
    e->chunkRamp = EnvInt("NNBVH_CHUNK_RAMP", e->chunkRamp);
    e->chunkRays = (uint64_t)EnvInt("NNBVH_CHUNK_RAYS", (int)e->chunkRays);
    e->chunkRaysSoa = (uint64_t)EnvInt("NNBVH_CHUNK_RAYS_SOA", (int)e->chunkRaysSoa);
    e->nSlots = std::min((int)nnbvh_engine::kMaxSlots, std::max(1, EnvInt("NNBVH_SLOTS", e->nSlots)));
    e->binMode = EnvInt("NNBVH_BIN", e->binMode);
    e->binThreshold = EnvInt("NNBVH_BIN_THRESHOLD", e->binThreshold);
    e->exactOnly = EnvInt("NNBVH_EXACT_ONLY", e->exactOnly);
    e->binCellBits = std::min(8, std::max(-1, EnvInt("NNBVH_BIN_CELL_BITS", e->binCellBits)));
    e->binDirBits = std::min(4, std::max(0, EnvInt("NNBVH_BIN_DIR_BITS", e->binDirBits)));
    e->binMinRays = (uint64_t)EnvInt("NNBVH_BIN_MIN_RAYS", (int)e->binMinRays);
    e->binMinNodes = (uint32_t)EnvInt("NNBVH_BIN_MIN_NODES", (int)e->binMinNodes);
    CU(cudaStreamCreateWithFlags(&e->ownStream, cudaStreamNonBlocking));
    e->stream = e->ownStream;
    CU(cudaEventCreate(&e->evA));
    CU(cudaEventCreate(&e->evB));
    CU(cudaEventCreate(&e->evC));
    CU(cudaEventCreate(&e->evD));
    CU(cudaMalloc(&e->scratch.cursor, kCursorWords * sizeof(unsigned long long)));
    for (Slot &s : e->slots) {
        CU(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
        CU(cudaEventCreate(&s.k0));
        CU(cudaEventCreate(&s.k1));
        CU(cudaMalloc(&s.scratch.cursor, kCursorWords * sizeof(unsigned long long)));
    }
    int rcSetup = NoAllocSetup(dev);
    if (rcSetup) return rcSetup;
    *out = owner.release();
    return NNBVH_OK;
}

extern "C" void nnbvh_destroy(nnbvh_engine *e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();
    e->store.reset();  // frees the device scene unless a clone still holds it
    if (e->dStats) cudaFree(e->dStats);
    e->scratch.Free();
    for (Slot &s : e->slots) {
        s.rays.Free(); s.soa.Free(); s.hits.Free(); s.occ.Free(); s.cnt.Free(); s.inner.Free();
        s.scratch.Free();
        if (s.done) cudaEventDestroy(s.done);
        if (s.k0) cudaEventDestroy(s.k0);
        if (s.k1) cudaEventDestroy(s.k1);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    if (e->evA) cudaEventDestroy(e->evA);
    if (e->evB) cudaEventDestroy(e->evB);
    if (e->evC) cudaEventDestroy(e->evC);
    if (e->evD) cudaEventDestroy(e->evD);
    if (e->ownStream) cudaStreamDestroy(e->ownStream);
    delete e;
}

extern "C" int nnbvh_set_stream(nnbvh_engine *e, void *cudaStream) {
    if (!e) return SetError(NNBVH_ERR_INVALID, "nnbvh_set_stream: engine is NULL");
    e->stream = cudaStream ? (cudaStream_t)cudaStream : e->ownStream;
    return NNBVH_OK;
}

extern "C" int nnbvh_clone(nnbvh_engine *src, nnbvh_engine **out) {
    if (!src || !out) return SetError(NNBVH_ERR_INVALID, "nnbvh_clone: NULL argument");
    nnbvh_options o{};
    o.device = src->device;
    o.kernel = src->kernel;
    o.smem_top_nodes = src->smemTopNodes;
    nnbvh_engine *e = nullptr;
    int rc = nnbvh_create(&o, &e);
    if (rc) return rc;
    e->store = src->store;  // the scene stays where it is
    e->sc = src->sc;
    e->haveScene = src->haveScene;
    e->boxesRegular = src->boxesRegular;
    e->hasSpheres = src->hasSpheres;
    e->hasInstances = src->hasInstances;
    e->hasAlpha = src->hasAlpha;
    e->general = src->general;
    e->instStepOK = src->instStepOK;
    e->generalStepOK = src->generalStepOK;
    *out = e;
    return NNBVH_OK;
}

extern "C" int nnbvh_last_kernel_ms(nnbvh_engine *e, float *ms) {
    if (!e || !ms) return SetError(NNBVH_ERR_INVALID, "nnbvh_last_kernel_ms: NULL argument");
    if (e->lastMs < 0.f) {  // device-pointer call: resolve the event pair now
        CU(cudaEventSynchronize(e->evB));
        float t = 0.f;
        CU(cudaEventElapsedTime(&t, e->evA, e->evB));
        e->lastMs = t;
    }
    *ms = e->lastMs;
    return NNBVH_OK;
}

extern "C" int nnbvh_last_traversal_ms(nnbvh_engine *e, float *ms) {
    if (!e || !ms) return SetError(NNBVH_ERR_INVALID, "nnbvh_last_traversal_ms: NULL argument");
    if (!e->haveTraversalEvents)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_last_traversal_ms: no device-pointer batch has run "
                                           "through the StepKernel path yet");
    CU(cudaSetDevice(e->device));
    CU(cudaEventSynchronize(e->evD));
    CU(cudaEventElapsedTime(ms, e->evC, e->evD));
    return NNBVH_OK;
}

extern "C" int nnbvh_debug_step_stats(nnbvh_engine *e, uint64_t *out8) {
    if (!e || !out8) return SetError(NNBVH_ERR_INVALID, "nnbvh_debug_step_stats: NULL argument");
    if (!e->dStats) return SetError(NNBVH_ERR_INVALID, "nnbvh_debug_step_stats: NNBVH_STEP_STATS was not set");
    CU(cudaSetDevice(e->device));
    CU(cudaStreamSynchronize(e->stream));
    CU(cudaMemcpy(out8, e->dStats, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    CU(cudaMemset(e->dStats, 0, 8 * sizeof(uint64_t)));
    return NNBVH_OK;
}

extern "C" int nnbvh_launch_count(nnbvh_engine *e, uint64_t *launches) {
    if (!e || !launches) return SetError(NNBVH_ERR_INVALID, "nnbvh_launch_count: NULL argument");
    *launches = e->launches;
    return NNBVH_OK;
}

// ---------------------------------------------------------------------------
// Scene upload: reference layout -> device layout
// ---------------------------------------------------------------------------
namespace nnbvh {

static bool IsDevicePtr(const void *p) {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// in9: nPrims*9 floats (ordered); last: 1 where the primitive ends its leaf.
// sphereOf / instanceOf (each optional): >= 0 marks a sphere / instance slot, the value is its index in the
// sphere / instance table.
__global__ void PackTrisKernel(const float *__restrict__ in9, const uint8_t *__restrict__ last,
                               const int32_t *__restrict__ sphereOf, const int32_t *__restrict__ instanceOf,
                               uint32_t nPrims, float4 *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPrims) return;
    const int32_t sph = sphereOf ? sphereOf[i] : -1, inst = instanceOf ? instanceOf[i] : -1;
    if (sph >= 0 || inst >= 0) {
        const uint32_t flag = inst >= 0 ? kPrimInstance : kPrimSphere;
        out[3 * (size_t)i + 0] = make_float4(__uint_as_float((uint32_t)(inst >= 0 ? inst : sph)), 0.f, 0.f,
                                             __uint_as_float(flag | (last[i] ? kTriLast : 0u)));
        out[3 * (size_t)i + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
        out[3 * (size_t)i + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
        return;
    }
    const float *p = in9 + 9 * (size_t)i;
    float v[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) v[k] = p[k];
    uint32_t flags = 0;
    if (TriangleDegenerate(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8]))
        flags |= kTriDegenerate;
    if (last[i]) flags |= kTriLast;
    out[3 * (size_t)i + 0] = make_float4(v[0], v[1], v[2], __uint_as_float(flags));
    out[3 * (size_t)i + 1] = make_float4(v[3], v[4], v[5], 0.f);
    out[3 * (size_t)i + 2] = make_float4(v[6], v[7], v[8], 0.f);
}

// Validates the tree and derives the wide layout.  Returns NNBVH_OK or an error.
static int BuildWideLayout(const std::vector<nnbvh_node> &nodes, uint32_t nPrims,
                           uint32_t topCap, std::vector<WideNode> &wide,
                           std::vector<uint8_t> &lastFlag, int32_t *rootRef,
                           uint32_t *topNodes) {
    const uint32_t M = (uint32_t)nodes.size();
    lastFlag.assign(nPrims, 0);
    std::vector<uint8_t> covered(nPrims, 0);
    for (uint32_t i = 0; i < M; ++i) {
        const nnbvh_node &nd = nodes[i];
        if (nd.nPrimitives > 0) {
            if (nd.offset < 0 || (uint64_t)nd.offset + nd.nPrimitives > nPrims)
                return SetError(NNBVH_ERR_INVALID, "tree: leaf %u covers primitives [%d,%d) of %u",
                                i, nd.offset, nd.offset + nd.nPrimitives, nPrims);
            lastFlag[nd.offset + nd.nPrimitives - 1] = 1;
            for (uint32_t k = 0; k < nd.nPrimitives; ++k) covered[nd.offset + k] = 1;
        } else {
            if (nd.axis > 2 || i + 1 >= M || nd.offset <= (int32_t)i + 1 ||
                (uint32_t)nd.offset >= M)
                return SetError(NNBVH_ERR_INVALID, "tree: interior node %u has axis %u, second "
                                "child %d (of %u nodes)", i, nd.axis, nd.offset, M);
        }
    }
    uint32_t depth = 0;
    int rc = nnbvh_tree_stack_depth(nodes.data(), M, &depth);
    if (rc) return rc;
    if (depth > (uint32_t)kMaxStack)
        return SetError(NNBVH_ERR_DEPTH, "tree: traversal needs %u stack entries; the reference's "
                        "nodesToVisit[64] (cpu/aggregates.cpp:538) would overflow", depth);
    // Number the interior nodes: BFS for the top (shared-memory resident,
    // contiguous), pre-order for the rest (keeps subtrees local in memory).
    std::vector<int32_t> wideOf(M, -1);
    uint32_t nWide = 0;
    if (nodes[0].nPrimitives == 0) {
        std::deque<uint32_t> q{0u};
        while (!q.empty() && nWide < topCap) {
            uint32_t i = q.front();
            q.pop_front();
            wideOf[i] = (int32_t)nWide++;
            uint32_t c0 = i + 1, c1 = (uint32_t)nodes[i].offset;
            if (nodes[c0].nPrimitives == 0) q.push_back(c0);
            if (nodes[c1].nPrimitives == 0) q.push_back(c1);
        }
        *topNodes = nWide;
        for (uint32_t i = 0; i < M; ++i)  // array order IS pre-order
            if (nodes[i].nPrimitives == 0 && wideOf[i] < 0) wideOf[i] = (int32_t)nWide++;
    } else {
        *topNodes = 0;
    }
    wide.assign(nWide, WideNode{});
    auto refOf = [&](uint32_t i) -> int32_t {
        return nodes[i].nPrimitives == 0 ? wideOf[i]
                                         : (int32_t)(0x80000000u | (uint32_t)nodes[i].offset);
    };
    for (uint32_t i = 0; i < M; ++i) {
        if (nodes[i].nPrimitives != 0) continue;
        WideNode &w = wide[wideOf[i]];
        const nnbvh_node &a = nodes[i + 1], &b = nodes[nodes[i].offset];
        for (int k = 0; k < 3; ++k) {
            w.lo0[k] = a.bounds[k];
            w.hi0[k] = a.bounds[3 + k];
            w.lo1[k] = b.bounds[k];
            w.hi1[k] = b.bounds[3 + k];
        }
        w.ref0 = refOf(i + 1);
        w.ref1 = refOf((uint32_t)nodes[i].offset);
        w.axis = nodes[i].axis;
        w.axisBit = 1 << nodes[i].axis;
    }
    *rootRef = refOf(0);
    return NNBVH_OK;
}

}  // namespace nnbvh

// An upload into an engine that shares its scene with clones gets a fresh store: the clones keep theirs.
static void DetachStore(nnbvh_engine *e) {
    if (e && e->store.use_count() > 1) e->store = std::make_shared<SceneStore>();
}

// Sphere::Sphere (shapes.h:117-128): the clamped zMin / zMax the intersection routine uses.
static int MakeSphereDev(const nnbvh_sphere &s, uint32_t index, SphereDev *out) {
    for (int k = 0; k < 12; ++k)
        if (!std::isfinite(s.objectFromRender[k]) || !std::isfinite(s.renderFromObject[k]))
            return SetError(NNBVH_ERR_INVALID, "sphere %u: non-finite transform", index);
    if (!(s.radius > 0.f) || !std::isfinite(s.radius))
        return SetError(NNBVH_ERR_INVALID, "sphere %u: radius %g", index, (double)s.radius);
    auto clamp = [](float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); };
    for (int k = 0; k < 12; ++k) out->m[k] = s.objectFromRender[k];
    out->radius = s.radius;
    out->zMin = clamp(std::min(s.zMin, s.zMax), -s.radius, s.radius);
    out->zMax = clamp(std::max(s.zMin, s.zMax), -s.radius, s.radius);
    out->phiMax = (3.14159265358979323846f / 180) * clamp(s.phiMax, 0.f, 360.f);  // Radians(), util/math.h:261-263
    return NNBVH_OK;
}

// The objects of an instanced scene as StepKernel<..., INST> wants them: WideNodes and triangles that UploadScene
// appends to the top level's arrays (refs are relative to the start of the object level until it rebases them).
struct ObjectLevel {
    std::vector<WideNode> wide;
    std::vector<float> tris9;
    std::vector<uint8_t> last;
    std::vector<uint32_t> wideStart, primStart;  // per object: where it starts in `wide` / in the triangles
    std::vector<int32_t> rootRef;                // per object, relative to the object level
    std::vector<float> rootBox;                  // 6 floats per object
    std::vector<uint32_t> objectOf;              // per instance
    uint32_t maxDepth = 0;                       // deepest traversal stack any object needs
    bool regular = true;                         // every box finite with lo <= hi
};

static int UploadScene(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                       const float *orderedTriVerts9, uint32_t nPrims, const int32_t *sphereOf,
                       const nnbvh_sphere *spheres, uint32_t nSpheres,
                       const int32_t *instanceOf = nullptr, uint32_t nInstances = 0,
                       ObjectLevel *objs = nullptr);

extern "C" int nnbvh_upload_scene(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                                  const float *orderedTriVerts9, uint32_t nPrims) {
    DetachStore(e);
    return UploadScene(e, nodes, nNodes, orderedTriVerts9, nPrims, nullptr, nullptr, 0);
}

extern "C" int nnbvh_upload_scene_alpha(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                                        const float *orderedTriVerts9, const float *orderedAlpha, uint32_t nPrims) {
    if (nPrims && !orderedAlpha) return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_alpha: NULL alpha array");
    if (IsDevicePtr(orderedAlpha))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_alpha: the alpha array is a host array");
    for (uint32_t i = 0; i < nPrims; ++i)
        if (std::isnan(orderedAlpha[i]))
            return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_alpha: alpha of primitive %u is NaN", i);
    DetachStore(e);
    int rc = UploadScene(e, nodes, nNodes, orderedTriVerts9, nPrims, nullptr, nullptr, 0);
    if (rc || nNodes == 0 || nPrims == 0) return rc;
    rc = e->store->dAlpha.Reserve(nPrims);
    if (rc) return rc;
    cudaStream_t st = e->stream;
    CU(cudaMemcpyAsync(e->store->dAlpha.p, orderedAlpha, sizeof(float) * (size_t)nPrims, cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));
    e->sc.alpha = e->store->dAlpha.p;
    e->hasAlpha = true;
    return NNBVH_OK;
}

extern "C" int nnbvh_upload_scene_mixed(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                                        const float *orderedTriVerts9, const int32_t *orderedSphereIndex,
                                        uint32_t nPrims, const nnbvh_sphere *spheres, uint32_t nSpheres) {
    if (nSpheres && (!spheres || !orderedSphereIndex))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_mixed: NULL sphere table / index");
    if (IsDevicePtr(orderedSphereIndex) || IsDevicePtr(spheres))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_mixed: the sphere table and index are host arrays");
    DetachStore(e);
    return UploadScene(e, nodes, nNodes, orderedTriVerts9, nPrims, nSpheres ? orderedSphereIndex : nullptr,
                       spheres, nSpheres);
}

static int UploadScene(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                       const float *orderedTriVerts9, uint32_t nPrims, const int32_t *sphereOf,
                       const nnbvh_sphere *spheres, uint32_t nSpheres, const int32_t *instanceOf,
                       uint32_t nInstances, ObjectLevel *objs) {
    if (!e) return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene: engine is NULL");
    CU(cudaSetDevice(e->device));
    e->haveScene = false;
    e->hasSpheres = false;
    e->hasInstances = false;
    e->hasAlpha = false;
    e->general = false;
    e->instStepOK = false;
    e->generalStepOK = false;
    if (nNodes == 0) {  // `nodes == nullptr` in the reference: every ray misses
        e->sc = SceneDev{};
        e->haveScene = true;
        return NNBVH_OK;
    }
    if (!nodes || !orderedTriVerts9 || nPrims == 0)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene: NULL nodes/triangles");
    if (nPrims >= 0x80000000u)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene: too many primitives");
    cudaStream_t st = e->stream;
    std::vector<nnbvh_node> h(nNodes);
    CU(cudaMemcpyAsync(h.data(), nodes, sizeof(nnbvh_node) * (size_t)nNodes, cudaMemcpyDefault, st));
    CU(cudaStreamSynchronize(st));
    e->boxesRegular = true;
    for (const nnbvh_node &nd : h)
        for (int k = 0; k < 3; ++k) {
            float lo = nd.bounds[k], hi = nd.bounds[3 + k];
            if (!(std::isfinite(lo) && std::isfinite(hi) && lo <= hi)) e->boxesRegular = false;
        }
    std::vector<WideNode> wide;
    std::vector<uint8_t> last;
    int32_t rootRef = 0;
    uint32_t topNodes = 0;
    const uint32_t topCap = 2048;  // 128 KiB of WideNodes at most in shared memory
    int rc = BuildWideLayout(h, nPrims, topCap, wide, last, &rootRef, &topNodes);
    if (rc) return rc;
    // Instanced scenes on the StepKernel: the objects' WideNodes follow the top level's, their triangles follow
    // the top level's triangles; refs become absolute.
    const uint32_t nWideTop = (uint32_t)wide.size();
    uint32_t nObjPrims = 0;
    bool stepInstances = false;
    if (objs) {
        uint32_t depthTop = 0;
        rc = nnbvh_tree_stack_depth(h.data(), nNodes, &depthTop);
        if (rc) return rc;
        nObjPrims = (uint32_t)objs->last.size();
        stepInstances = depthTop + 1 + objs->maxDepth <= (uint32_t)kMaxStack &&
                        (uint64_t)nPrims + nObjPrims < 0x80000000ull;
        if (stepInstances) {
            if (!objs->regular) e->boxesRegular = false;
            wide.reserve(nWideTop + objs->wide.size());
            for (WideNode w : objs->wide) {
                w.ref0 = w.ref0 >= 0 ? w.ref0 + (int32_t)nWideTop : (int32_t)(0x80000000u | (((uint32_t)w.ref0 & 0x7fffffffu) + nPrims));
                w.ref1 = w.ref1 >= 0 ? w.ref1 + (int32_t)nWideTop : (int32_t)(0x80000000u | (((uint32_t)w.ref1 & 0x7fffffffu) + nPrims));
                wide.push_back(w);
            }
        } else nObjPrims = 0;
    }
    rc = e->store->dNodes.Reserve(nNodes);
    if (rc) return rc;
    rc = e->store->dWide.Reserve(wide.size());
    if (rc) return rc;
    rc = e->store->dTris.Reserve(3 * ((size_t)nPrims + nObjPrims));
    if (rc) return rc;
    CU(cudaMemcpyAsync(e->store->dNodes.p, h.data(), sizeof(nnbvh_node) * (size_t)nNodes,
                       cudaMemcpyHostToDevice, st));
    if (!wide.empty())
        CU(cudaMemcpyAsync(e->store->dWide.p, wide.data(), sizeof(WideNode) * wide.size(),
                           cudaMemcpyHostToDevice, st));
    // triangles: stage raw floats + leaf-end flags, pack on the device (temporaries freed on every return path)
    DevTmp<float> dRaw;
    DevTmp<uint8_t> dLast;
    DevTmp<int32_t> dSphereOf;
    const float *src = orderedTriVerts9;
    if (!IsDevicePtr(orderedTriVerts9)) {
        CU(dRaw.Alloc(9 * (size_t)nPrims));
        CU(cudaMemcpyAsync(dRaw.p, orderedTriVerts9, sizeof(float) * 9 * (size_t)nPrims,
                           cudaMemcpyHostToDevice, st));
        src = dRaw.p;
    }
    CU(dLast.Alloc(nPrims));
    CU(cudaMemcpyAsync(dLast.p, last.data(), nPrims, cudaMemcpyHostToDevice, st));
    std::vector<SphereDev> sd;  // outlives the asynchronous copy below
    DevTmp<int32_t> dInstanceOf;
    if (sphereOf) {
        sd.resize(nSpheres);
        for (uint32_t i = 0; i < nSpheres; ++i) {
            rc = MakeSphereDev(spheres[i], i, &sd[i]);
            if (rc) break;
        }
        for (uint32_t i = 0; !rc && i < nPrims; ++i)
            if (sphereOf[i] >= (int32_t)nSpheres)
                rc = SetError(NNBVH_ERR_INVALID, "primitive %u refers to sphere %d of %u", i, sphereOf[i], nSpheres);
        if (!rc) rc = e->store->dSpheres.Reserve(nSpheres);
        if (rc) {
            cudaStreamSynchronize(st);  // the staged copies read host vectors of this frame
            return rc;
        }
        CU(cudaMemcpyAsync(e->store->dSpheres.p, sd.data(), sizeof(SphereDev) * nSpheres, cudaMemcpyHostToDevice, st));
        CU(dSphereOf.Alloc(nPrims));
        CU(cudaMemcpyAsync(dSphereOf.p, sphereOf, sizeof(int32_t) * (size_t)nPrims, cudaMemcpyHostToDevice, st));
    }
    if (instanceOf) {
        for (uint32_t i = 0; i < nPrims; ++i)
            if (instanceOf[i] >= (int32_t)nInstances || (instanceOf[i] >= 0 && sphereOf && sphereOf[i] >= 0)) {
                cudaStreamSynchronize(st);
                return SetError(NNBVH_ERR_INVALID, "primitive %u refers to instance %d of %u%s", i, instanceOf[i], nInstances,
                                instanceOf[i] < (int32_t)nInstances ? " and to a sphere as well" : "");
            }
        CU(dInstanceOf.Alloc(nPrims));
        CU(cudaMemcpyAsync(dInstanceOf.p, instanceOf, sizeof(int32_t) * (size_t)nPrims, cudaMemcpyHostToDevice, st));
    }
    PackTrisKernel<<<(nPrims + 255) / 256, 256, 0, st>>>(src, dLast.p, dSphereOf.p, dInstanceOf.p, nPrims, e->store->dTris.p);
    ++e->launches;
    DevTmp<float> dObjRaw;
    DevTmp<uint8_t> dObjLast;
    std::vector<InstanceStep> steps;  // outlives the asynchronous copy
    if (stepInstances) {
        CU(dObjRaw.Alloc(objs->tris9.size()));
        CU(dObjLast.Alloc(nObjPrims));
        CU(cudaMemcpyAsync(dObjRaw.p, objs->tris9.data(), sizeof(float) * objs->tris9.size(), cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(dObjLast.p, objs->last.data(), nObjPrims, cudaMemcpyHostToDevice, st));
        PackTrisKernel<<<(nObjPrims + 255) / 256, 256, 0, st>>>(dObjRaw.p, dObjLast.p, nullptr, nullptr, nObjPrims,
                                                                 e->store->dTris.p + 3 * (size_t)nPrims);
        ++e->launches;
        steps.resize(objs->objectOf.size());
        for (size_t k = 0; k < steps.size(); ++k) {
            const uint32_t j = objs->objectOf[k];
            for (int c = 0; c < 6; ++c) steps[k].rootBox[c] = objs->rootBox[6 * j + c];
            const int32_t r = objs->rootRef[j];
            steps[k].rootRef = r >= 0 ? r + (int32_t)nWideTop : (int32_t)(0x80000000u | (((uint32_t)r & 0x7fffffffu) + nPrims));
            steps[k].primBase = nPrims + objs->primStart[j];
        }
        rc = e->store->dInstStep.Reserve(steps.size());
        if (rc) {
            cudaStreamSynchronize(st);
            return rc;
        }
        CU(cudaMemcpyAsync(e->store->dInstStep.p, steps.data(), sizeof(InstanceStep) * steps.size(), cudaMemcpyHostToDevice, st));
    }
    cudaError_t launchErr = cudaGetLastError();
    cudaError_t syncErr = cudaStreamSynchronize(st);  // before the temporaries (and `h`, `wide`, `last`, `sd`) go away
    CU(launchErr);
    CU(syncErr);
    SceneDev sc{};
    sc.nodes = e->store->dNodes.p;
    sc.wide = e->store->dWide.p;
    sc.tris = e->store->dTris.p;
    sc.nNodes = nNodes;
    sc.nWide = (uint32_t)wide.size();
    sc.nPrims = nPrims;
    for (int k = 0; k < 6; ++k) sc.rootBox[k] = h[0].bounds[k];
    sc.rootRef = rootRef;
    sc.topNodes = topNodes;
    sc.spheres = sphereOf ? e->store->dSpheres.p : nullptr;
    sc.instStep = stepInstances ? e->store->dInstStep.p : nullptr;
    e->instStepOK = stepInstances;
    e->sc = sc;
    e->hasSpheres = sphereOf != nullptr;
    e->haveScene = true;
    return NNBVH_OK;
}

// TransformedPrimitive::Bounds() (cpu/primitive.h:94) = renderFromPrimitive(primitive.Bounds()):
// the eight transformed corners (util/transform.cpp:134-139, util/transform.h:310-319).
extern "C" int nnbvh_instance_bounds(const nnbvh_instance *in, const float *objectBounds6, float *bounds6) {
    if (!in || !objectBounds6 || !bounds6) return SetError(NNBVH_ERR_INVALID, "nnbvh_instance_bounds: NULL argument");
    float mn[3], mx[3];
    for (int k = 0; k < 3; ++k) {
        mn[k] = std::numeric_limits<float>::max();
        mx[k] = std::numeric_limits<float>::lowest();
    }
    const float *m = in->renderFromPrimitive;
    for (int c = 0; c < 8; ++c) {
        const float x = objectBounds6[(c & 1) ? 3 : 0], y = objectBounds6[(c & 2) ? 4 : 1], z = objectBounds6[(c & 4) ? 5 : 2];
        const float q[3] = {m[0] * x + m[1] * y + m[2] * z + m[3], m[4] * x + m[5] * y + m[6] * z + m[7],
                            m[8] * x + m[9] * y + m[10] * z + m[11]};
        for (int k = 0; k < 3; ++k) {
            mn[k] = std::min(mn[k], q[k]);
            mx[k] = std::max(mx[k], q[k]);
        }
    }
    for (int k = 0; k < 3; ++k) bounds6[k] = mn[k], bounds6[3 + k] = mx[k];
    return NNBVH_OK;
}

// The objects of an instanced scene: validate each tree, lay trees and triangles back to back on the device,
// fill the instance table.  `e`'s device is current.
static int UploadObjects(nnbvh_engine *e, const nnbvh_instance *instances, uint32_t nInstances,
                         const nnbvh_object *objects, uint32_t nObjects, ObjectLevel *lvl) {
    cudaStream_t st = e->stream;
    std::vector<nnbvh_node> allNodes;
    std::vector<float> allTris;
    std::vector<uint8_t> allLast;
    std::vector<uint32_t> nodeBase(nObjects), primBase(nObjects);
    for (uint32_t j = 0; j < nObjects; ++j) {
        const nnbvh_object &o = objects[j];
        if (!o.nodes || !o.orderedTriVerts9 || o.nNodes == 0 || o.nPrims == 0)
            return SetError(NNBVH_ERR_INVALID, "object %u is empty", j);
        std::vector<nnbvh_node> h(o.nodes, o.nodes + o.nNodes);
        std::vector<WideNode> wide;
        std::vector<uint8_t> last;
        int32_t rootRef;
        uint32_t topNodes;
        int rc = BuildWideLayout(h, o.nPrims, 0, wide, last, &rootRef, &topNodes);  // validation + leaf-end flags
        if (rc) return rc;
        nodeBase[j] = (uint32_t)allNodes.size();
        primBase[j] = (uint32_t)allLast.size();
        {   // the object in the StepKernel's layout: refs relative to the start of the object level
            const uint32_t w0 = (uint32_t)lvl->wide.size(), p0 = primBase[j];
            auto rel = [&](int32_t r) {
                return r >= 0 ? r + (int32_t)w0 : (int32_t)(0x80000000u | (((uint32_t)r & 0x7fffffffu) + p0));
            };
            lvl->wideStart.push_back(w0);
            lvl->primStart.push_back(p0);
            for (WideNode w : wide) {
                w.ref0 = rel(w.ref0);
                w.ref1 = rel(w.ref1);
                lvl->wide.push_back(w);
            }
            lvl->rootRef.push_back(rel(rootRef));
            for (int c = 0; c < 6; ++c) lvl->rootBox.push_back(h[0].bounds[c]);
            uint32_t depth = 0;
            rc = nnbvh_tree_stack_depth(h.data(), o.nNodes, &depth);
            if (rc) return rc;
            lvl->maxDepth = std::max(lvl->maxDepth, depth);
            for (const nnbvh_node &nd : h)
                for (int c = 0; c < 3; ++c) {
                    const float lo = nd.bounds[c], hi = nd.bounds[3 + c];
                    if (!(std::isfinite(lo) && std::isfinite(hi) && lo <= hi)) lvl->regular = false;
                }
        }
        allNodes.insert(allNodes.end(), h.begin(), h.end());
        allTris.insert(allTris.end(), o.orderedTriVerts9, o.orderedTriVerts9 + 9 * (size_t)o.nPrims);
        allLast.insert(allLast.end(), last.begin(), last.end());
    }
    std::vector<InstanceDev> id(nInstances);
    for (uint32_t k = 0; k < nInstances; ++k) {
        const nnbvh_instance &in = instances[k];
        if (in.object >= nObjects) return SetError(NNBVH_ERR_INVALID, "instance %u refers to object %u of %u", k, in.object, nObjects);
        for (int c = 0; c < 12; ++c) {
            if (!std::isfinite(in.primitiveFromRender[c]) || !std::isfinite(in.renderFromPrimitive[c]))
                return SetError(NNBVH_ERR_INVALID, "instance %u: non-finite transform", k);
            id[k].mInv[c] = in.primitiveFromRender[c];
        }
        lvl->objectOf.push_back(in.object);
        id[k].nodeBase = nodeBase[in.object];
        id[k].primBase = primBase[in.object];
        id[k].nNodes = objects[in.object].nNodes;
        id[k].pad = 0;
    }
    lvl->tris9 = allTris;
    lvl->last = allLast;
    const uint32_t nObjPrims = (uint32_t)allLast.size();
    int rc = e->store->dObjNodes.Reserve(allNodes.size());
    if (!rc) rc = e->store->dObjTris.Reserve(3 * (size_t)nObjPrims);
    if (!rc) rc = e->store->dInstances.Reserve(nInstances);
    if (rc) return rc;
    DevTmp<float> dRaw;
    DevTmp<uint8_t> dLast;
    cudaError_t ce = dRaw.Alloc(allTris.size());
    if (ce == cudaSuccess) ce = dLast.Alloc(nObjPrims);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(dRaw.p, allTris.data(), sizeof(float) * allTris.size(), cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(dLast.p, allLast.data(), nObjPrims, cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(e->store->dObjNodes.p, allNodes.data(), sizeof(nnbvh_node) * allNodes.size(), cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(e->store->dInstances.p, id.data(), sizeof(InstanceDev) * nInstances, cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) {
        PackTrisKernel<<<(nObjPrims + 255) / 256, 256, 0, st>>>(dRaw.p, dLast.p, nullptr, nullptr, nObjPrims, e->store->dObjTris.p);
        ++e->launches;
        ce = cudaGetLastError();
    }
    cudaError_t se = cudaStreamSynchronize(st);  // the copies read this frame's vectors; the temporaries go away
    CU(ce);
    CU(se);
    return NNBVH_OK;
}

extern "C" int nnbvh_upload_scene_instanced(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                                            const float *orderedTriVerts9, const int32_t *orderedInstanceIndex,
                                            uint32_t nPrims, const nnbvh_instance *instances, uint32_t nInstances,
                                            const nnbvh_object *objects, uint32_t nObjects) {
    if (!e) return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_instanced: engine is NULL");
    if (!instances || !objects || !orderedInstanceIndex || nInstances == 0 || nObjects == 0)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_instanced: NULL / empty instance or object table");
    CU(cudaSetDevice(e->device));
    DetachStore(e);
    ObjectLevel lvl;
    int rc = UploadObjects(e, instances, nInstances, objects, nObjects, &lvl);
    if (rc) return rc;
    // the top level: triangles and instance slots (+ the objects appended in the StepKernel's layout)
    rc = UploadScene(e, nodes, nNodes, orderedTriVerts9, nPrims, nullptr, nullptr, 0, orderedInstanceIndex, nInstances, &lvl);
    if (rc) return rc;
    e->sc.instances = e->store->dInstances.p;
    e->sc.objNodes = e->store->dObjNodes.p;
    e->sc.objTris = e->store->dObjTris.p;
    e->hasInstances = true;
    return NNBVH_OK;
}

// One primitive list holding any mix of triangles (some with a constant alpha), spheres and instances
// (cpu/primitive.cpp:50-130 behind one BVHAggregate, as scene.cpp:1352-1591 builds it).  A description that
// uses at most one of the three extras takes that extra's own path (StepKernel for spheres, ...); a real
// mix runs the general instantiation of BasicKernel (one thread per ray, nested traversal, alpha re-trace).
extern "C" int nnbvh_upload_scene_general(nnbvh_engine *e, const nnbvh_scene_desc *d) {
    if (!e || !d) return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_general: NULL argument");
    const bool sph = d->nSpheres > 0, inst = d->nInstances > 0;
    bool alpha = false;
    if (d->orderedAlpha) {
        if (IsDevicePtr(d->orderedAlpha))
            return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_general: the alpha array is a host array");
        for (uint32_t i = 0; i < d->nPrims; ++i) {
            if (std::isnan(d->orderedAlpha[i]))
                return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_general: alpha of primitive %u is NaN", i);
            alpha |= d->orderedAlpha[i] < 1;
        }
    }
    if (sph && (!d->spheres || !d->orderedSphereIndex))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_general: NULL sphere table / index");
    if (inst && (!d->instances || !d->objects || !d->orderedInstanceIndex || d->nObjects == 0))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_general: NULL / empty instance or object table");
    if ((sph && (IsDevicePtr(d->orderedSphereIndex) || IsDevicePtr(d->spheres))) ||
        (inst && (IsDevicePtr(d->orderedInstanceIndex) || IsDevicePtr(d->instances))))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_upload_scene_general: tables and index arrays are host arrays");
    const int kinds = (int)sph + (int)inst + (int)alpha;
    if (kinds <= 1) {  // the specialised paths
        if (sph) return nnbvh_upload_scene_mixed(e, d->nodes, d->nNodes, d->orderedTriVerts9, d->orderedSphereIndex,
                                                 d->nPrims, d->spheres, d->nSpheres);
        if (inst) return nnbvh_upload_scene_instanced(e, d->nodes, d->nNodes, d->orderedTriVerts9, d->orderedInstanceIndex,
                                                      d->nPrims, d->instances, d->nInstances, d->objects, d->nObjects);
        if (alpha) return nnbvh_upload_scene_alpha(e, d->nodes, d->nNodes, d->orderedTriVerts9, d->orderedAlpha, d->nPrims);
        return nnbvh_upload_scene(e, d->nodes, d->nNodes, d->orderedTriVerts9, d->nPrims);
    }
    CU(cudaSetDevice(e->device));
    DetachStore(e);
    int rc = NNBVH_OK;
    ObjectLevel lvl;
    if (inst) rc = UploadObjects(e, d->instances, d->nInstances, d->objects, d->nObjects, &lvl);
    if (rc) return rc;
    rc = UploadScene(e, d->nodes, d->nNodes, d->orderedTriVerts9, d->nPrims, sph ? d->orderedSphereIndex : nullptr,
                     d->spheres, d->nSpheres, inst ? d->orderedInstanceIndex : nullptr, d->nInstances,
                     inst ? &lvl : nullptr);
    if (rc || d->nNodes == 0) return rc;
    // alpha: always present on this path (1 = none), so the kernel needs no NULL test
    std::vector<float> a(d->nPrims, 1.f);
    if (alpha) a.assign(d->orderedAlpha, d->orderedAlpha + d->nPrims);
    rc = e->store->dAlpha.Reserve(d->nPrims);
    if (rc) return rc;
    CU(cudaMemcpyAsync(e->store->dAlpha.p, a.data(), sizeof(float) * (size_t)d->nPrims, cudaMemcpyHostToDevice, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    e->sc.alpha = e->store->dAlpha.p;
    if (inst) {
        e->sc.instances = e->store->dInstances.p;
        e->sc.objNodes = e->store->dObjNodes.p;
        e->sc.objTris = e->store->dObjTris.p;
    }
    e->hasSpheres = false;  // the flags of the specialised paths stay off: `general` selects the kernel
    e->general = true;
    e->generalStepOK = !inst || e->instStepOK;  // the objects fit the StepKernel's combined stack
    return NNBVH_OK;
}

// ---------------------------------------------------------------------------
// Launch
// ---------------------------------------------------------------------------
namespace nnbvh {

struct Outputs {
    nnbvh_hit *hits = nullptr;
    uint8_t *occ = nullptr;
    uint32_t *nodesVisited = nullptr, *triTests = nullptr;
    uint32_t *inner = nullptr;  // instanced scenes, closest hit: primitive inside the object (optional)
};

template <bool ANY, int MINB>
static int LaunchWideT(nnbvh_engine *e, cudaStream_t st, const WideParams &p, uint32_t smemBytes) {
    static int configuredFor[16] = {};  // per device: max dynamic smem already opted in
    if (e->device < 16 && configuredFor[e->device] < (int)smemBytes + 1) {
        CU(cudaFuncSetAttribute(WideKernel<ANY, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                e->maxSmemOptin - 1024));
        configuredFor[e->device] = e->maxSmemOptin;
    }
    int grid = e->numSMs * MINB;
    WideKernel<ANY, MINB><<<grid, 256, smemBytes, st>>>(p);
    return NNBVH_OK;
}

template <bool ANY>
static int LaunchWide(nnbvh_engine *e, cudaStream_t st, const WideParams &p, uint32_t smemBytes) {
    switch (e->wideMinB) {
    case 1: return LaunchWideT<ANY, 1>(e, st, p, smemBytes);
    case 3: return LaunchWideT<ANY, 3>(e, st, p, smemBytes);
    case 2: return LaunchWideT<ANY, 2>(e, st, p, smemBytes);
    default: return LaunchWideT<ANY, 4>(e, st, p, smemBytes);
    }
}

// Grid of the persistent StepKernel: numSMs x MINB CTAs, fewer for small batches so that every lane still gets
// `raysPerLane` rays to refill from (a warp that holds one ray per lane for its whole life has nothing to
// replace finished rays with: NNBVH_GRID_RAYS_PER_LANE, 0 = always the full grid).
static unsigned StepGrid(const nnbvh_engine *e, int minB, uint64_t n) {
    const unsigned full = (unsigned)(e->numSMs * minB);
    if (e->gridRaysPerLane <= 0) return full;
    const uint64_t want = (n + 256ull * (uint64_t)e->gridRaysPerLane - 1) / (256ull * (uint64_t)e->gridRaysPerLane);
    return (unsigned)std::min<uint64_t>(full, std::max<uint64_t>(want, (uint64_t)e->numSMs));
}

template <bool ANY, bool EXACT>
static void LaunchStep(nnbvh_engine *e, cudaStream_t st, const StepParams &p) {
    const uint64_t n = p.nPtr ? 0 : p.n;  // the exact launch over the deferred list does not know its size on the host
    if (e->general) {  // every primitive kind in one list: spheres + instance slots + constant alpha, 3 CTAs/SM
        StepKernel<ANY, EXACT, 3, false, true, true, true><<<p.nPtr ? e->numSMs * 3 : StepGrid(e, 3, n), 256, 0, st>>>(p);
        return;
    }
    if (e->hasSpheres) {  // sphere leaves: one instantiation, 4 CTAs/SM (the sphere test wants registers)
        StepKernel<ANY, EXACT, 4, false, true><<<p.nPtr ? e->numSMs * 4 : StepGrid(e, 4, n), 256, 0, st>>>(p);
        return;
    }
    if (e->hasAlpha) {  // constant-alpha triangles: the alpha test inside the leaf phase, 4 CTAs/SM
        StepKernel<ANY, EXACT, 4, false, false, false, true><<<p.nPtr ? e->numSMs * 4 : StepGrid(e, 4, n), 256, 0, st>>>(p);
        return;
    }
    if (e->hasInstances) {  // instance slots: one instantiation, 4 CTAs/SM (the lane carries the instance state)
        StepKernel<ANY, EXACT, 4, false, false, true><<<p.nPtr ? e->numSMs * 4 : StepGrid(e, 4, n), 256, 0, st>>>(p);
        return;
    }
    if (p.stats && !EXACT) {  // NNBVH_STEP_STATS=1: lane-occupancy counters (diagnostic build of the kernel)
        StepKernel<ANY, false, 5, true><<<e->numSMs * 5, 256, 0, st>>>(p);
        return;
    }
    const int minB = (e->wideMinB >= 1 && e->wideMinB <= 6) ? e->wideMinB : 5;
    const unsigned grid = p.nPtr ? (unsigned)(e->numSMs * minB) : StepGrid(e, minB, n);
    switch (minB) {
    case 1: StepKernel<ANY, EXACT, 1><<<grid, 256, 0, st>>>(p); break;
    case 2: StepKernel<ANY, EXACT, 2><<<grid, 256, 0, st>>>(p); break;
    case 3: StepKernel<ANY, EXACT, 3><<<grid, 256, 0, st>>>(p); break;
    case 4: StepKernel<ANY, EXACT, 4><<<grid, 256, 0, st>>>(p); break;
    case 6: StepKernel<ANY, EXACT, 6><<<grid, 256, 0, st>>>(p); break;
    default: {
        // The kernel has no block-level state (warps are independent): at its 48 registers an SM holds 1365 threads,
        // i.e. 5 CTAs of 256 (40 warps) or 7 of 192 (42 warps).  NNBVH_STEP_BLOCK selects the CTA size.
        const unsigned block = (unsigned)e->stepBlock;
        if (block == 256 || p.nPtr || e->gridRaysPerLane > 0) {
            StepKernel<ANY, EXACT, 5><<<grid, 256, 0, st>>>(p);
        } else {
            const unsigned perSM = std::min(32u, 1344u / block);
            StepKernel<ANY, EXACT, 5><<<(unsigned)e->numSMs * perSM, block, 0, st>>>(p);
        }
        break;
    }
    }
}


// The ray-ordering pre-pass (binning.cuh): a permutation of the batch that keeps warps coherent, or nothing
// when the batch is too small / the mode says so.  cursor: the batch's zeroed scratch words (3, 4: sample
// statistics, 5: the device-side "sorted" flag).  *perm stays NULL when no sort was enqueued; *permOn is the
// device flag the kernels read (mode 2 decides per batch on the device).
static int OrderRays(nnbvh_engine *e, cudaStream_t st, const RayInput &in, uint64_t n, Scratch &scr,
                     unsigned long long *cursor, bool deviceDecides, uint32_t minNodes, const uint32_t **permOut,
                     const uint32_t **permOnOut) {
    const uint32_t *perm = nullptr, *permOn = nullptr;
    *permOut = *permOnOut = nullptr;
    if (!((e->binMode == 1 || (e->binMode == 2 && deviceDecides && e->sc.nNodes >= minNodes)) &&
          n >= e->binMinRays && n < 0xffffffffull && e->sc.nNodes))
        return NNBVH_OK;
    {
    BinParams bp;
    const size_t sceneBytes = sizeof(WideNode) * (size_t)e->sc.nWide + 48 * (size_t)e->sc.nPrims;
    bp.cellBits = (uint32_t)(e->binCellBits >= 0 ? e->binCellBits : (sceneBytes < ((size_t)96 << 20) ? 5 : 7));
    bp.dirBits = (uint32_t)e->binDirBits;
    for (int k = 0; k < 3; ++k) {
        float ext = e->sc.rootBox[3 + k] - e->sc.rootBox[k];
        bp.lo[k] = e->sc.rootBox[k];
        bp.scale[k] = ext > 0.f ? (float)(1u << bp.cellBits) / ext : 0.f;
    }
    const int keyBits = (int)(3 * bp.cellBits + 2 * bp.dirBits);
    const uint32_t nSortBlk = (uint32_t)((n + kSortTile - 1) / kSortTile);
    constexpr int kBinRadixBits = 7;  // 21-bit default key = 3 passes; runs of 16 pairs per digit
    const uint32_t nHist = (1u << kBinRadixBits) * nSortBlk, nTiles = (nHist + kScanTile - 1) / kScanTile;
    int rc = scr.keys.Reserve(n);
    if (!rc) rc = scr.keys2.Reserve(n);
    if (!rc) rc = scr.perm.Reserve(n);
    if (!rc) rc = scr.perm2.Reserve(n);
    if (!rc) rc = scr.hist.Reserve(nHist);
    if (!rc) rc = scr.tileSums.Reserve(nTiles);
    if (rc) return rc;
    uint32_t *flag = reinterpret_cast<uint32_t *>(cursor + 5);
    if (!scr.hFlag) {
        CU(cudaHostAlloc(&scr.hFlag, sizeof(uint32_t), cudaHostAllocDefault));
        *scr.hFlag = 1u;
    }
    const uint32_t lastFlag = *const_cast<volatile uint32_t *>(scr.hFlag);
    scr.coherentRun = lastFlag ? 0 : std::min(scr.coherentRun + 1, 1 << 20);
    const bool enqueueSort = e->binMode == 1 || scr.coherentRun < 2;
    BinSampleKernel<<<kBinSamples, 32, 0, st>>>(bp, in, n, cursor + 3);
    BinDecideKernel<<<1, 1, 0, st>>>(cursor + 3, (uint32_t)e->binThreshold,
                                     e->binMode == 1 ? 1u : 0u, flag);
    CU(cudaMemcpyAsync(scr.hFlag, flag, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    e->launches += 2;
    uint32_t *kIn = scr.keys.p, *kOut = scr.keys2.p, *vIn = nullptr, *vOut = scr.perm.p;
    if (enqueueSort) {
        BinKeyKernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(bp, in, n, scr.keys.p, flag);
        ++e->launches;
    }
    for (int shift = 0; enqueueSort && shift < keyBits; shift += kBinRadixBits) {
        RadixCountKernel<kBinRadixBits><<<nSortBlk, 256, 0, st>>>(kIn, (uint32_t)n, shift, scr.hist.p, flag);
        BinTileSumKernel<<<nTiles, 1024, 0, st>>>(scr.hist.p, nHist, scr.tileSums.p, flag);
        BinTileScanKernel<<<nTiles, 1024, 0, st>>>(scr.hist.p, nHist, scr.tileSums.p, flag);
        const bool lastPass = shift + kBinRadixBits >= keyBits;  // only the permutation is read afterwards
        RadixScatterKernel<kBinRadixBits><<<nSortBlk, 256, 0, st>>>(kIn, vIn, (uint32_t)n, shift, scr.hist.p,
                                                                   lastPass ? nullptr : kOut, vOut, flag);
        e->launches += 4;
        std::swap(kIn, kOut);
        uint32_t *done = vOut;
        vOut = (done == scr.perm.p) ? scr.perm2.p : scr.perm.p;
        vIn = done;
    }
    CU(cudaGetLastError());
    if (enqueueSort) {
        perm = vIn;
        permOn = flag;
    }
    }
    *permOut = perm;
    *permOnOut = permOn;
    return NNBVH_OK;
}

// All pointers are device pointers here.
static int LaunchTraversal(nnbvh_engine *e, cudaStream_t st, bool any, bool counters,
                           const RayInput &in, uint64_t n, const Outputs &o,
                           Scratch &scr) {
    if (n == 0) return NNBVH_OK;
    if (e->l2PersistBytes && e->l2MaxWindow && e->sc.nWide) {  // experiment: keep the wide nodes in the L2's set-aside
        cudaStreamAttrValue v{};
        const size_t bytes = std::min(sizeof(WideNode) * (size_t)e->sc.nWide, e->l2MaxWindow);
        v.accessPolicyWindow.base_ptr = const_cast<WideNode *>(e->sc.wide);
        v.accessPolicyWindow.num_bytes = bytes;
        v.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)e->l2PersistBytes / (double)bytes);
        v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        CU(cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v));
    }
    // One thread per ray (general / alpha / instanced scenes): the same ray order helps there too -- the lanes of a
    // warp then walk neighbouring rays -- so those kernels pull their ray through the permutation as well.
    RayInput inO = in;
    const bool generalStepEarly = e->general && e->generalStepOK && e->instStep && !counters &&
                                  e->kernel != NNBVH_KERNEL_BASIC && e->kernel != NNBVH_KERNEL_WIDE;
    const bool alphaStepEarly = e->hasAlpha && e->instStep && !counters && e->kernel != NNBVH_KERNEL_BASIC &&
                                e->kernel != NNBVH_KERNEL_WIDE;
    if (((e->general && !generalStepEarly) || (e->hasAlpha && !alphaStepEarly) ||
         (e->hasInstances && !(e->instStepOK && e->instStep && e->kernel != NNBVH_KERNEL_WIDE))) &&
        !counters && e->kernel != NNBVH_KERNEL_BASIC) {
        CU(cudaMemsetAsync(scr.cursor, 0, kCursorWords * sizeof(unsigned long long), st));
        // (no node-count floor here: an instanced scene's top level is small, its objects are not)
        int rc = OrderRays(e, st, in, n, scr, scr.cursor, true, 0u, &inO.perm, &inO.permOn);
        if (rc) return rc;
    }
    const bool generalStep = e->general && e->generalStepOK && e->instStep && !counters &&
                             e->kernel != NNBVH_KERNEL_BASIC && e->kernel != NNBVH_KERNEL_WIDE;
    if (e->general && !generalStep) {  // triangles + spheres + instances + alpha in one list: BasicKernel's general instantiation
        if (counters)
            return SetError(NNBVH_ERR_INVALID, "per-ray counters are not available for general (mixed-kind) scenes");
        if (n > (uint64_t)0x7fffffff * 128)
            return SetError(NNBVH_ERR_INVALID, "batch too large for one launch");
        const unsigned grid = (unsigned)((n + 127) / 128);
        if (any) BasicKernel<true, false, true, true, true><<<grid, 128, 0, st>>>(e->sc, inO, n, o.hits, o.occ, nullptr, nullptr, o.inner);
        else BasicKernel<false, false, true, true, true><<<grid, 128, 0, st>>>(e->sc, inO, n, o.hits, o.occ, nullptr, nullptr, o.inner);
        ++e->launches;
        CU(cudaGetLastError());
        return NNBVH_OK;
    }
    const bool alphaStep = e->hasAlpha && e->instStep && !counters && e->kernel != NNBVH_KERNEL_BASIC &&
                           e->kernel != NNBVH_KERNEL_WIDE;
    if (e->hasAlpha && !alphaStep) {  // alpha-tested primitives: one thread per ray (alpha.cuh)
        if (counters)
            return SetError(NNBVH_ERR_INVALID, "per-ray counters are not available for scenes with alpha primitives");
        if (n > (uint64_t)0x7fffffff * 128)
            return SetError(NNBVH_ERR_INVALID, "batch too large for one launch");
        const unsigned grid = (unsigned)((n + 127) / 128);
        if (any) AlphaKernel<true><<<grid, 128, 0, st>>>(e->sc, e->store->dAlpha.p, inO, n, o.hits, o.occ);
        else AlphaKernel<false><<<grid, 128, 0, st>>>(e->sc, e->store->dAlpha.p, inO, n, o.hits, o.occ);
        ++e->launches;
        CU(cudaGetLastError());
        return NNBVH_OK;
    }
    // instanced scenes: StepKernel<..., INST> when the combined stack depth fits (NNBVH_INST_STEP=0: BasicKernel);
    // counters and deeper scenes run the one-thread-per-ray kernel with its per-level stacks
    const bool instStep = generalStep || (e->hasInstances && e->instStepOK && e->instStep && e->kernel != NNBVH_KERNEL_WIDE);
    const bool basic = counters || e->kernel == NNBVH_KERNEL_BASIC || (e->hasInstances && !instStep);
    if (basic) {
        if (n > (uint64_t)0x7fffffff * 128)
            return SetError(NNBVH_ERR_INVALID, "batch too large for one launch");
        unsigned grid = (unsigned)((n + 127) / 128);
        if (e->hasInstances) {
            uint32_t *nv = counters ? o.nodesVisited : nullptr, *tt = counters ? o.triTests : nullptr;
            if (counters) {
                if (any) BasicKernel<true, true, false, true><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, nv, tt, o.inner);
                else BasicKernel<false, true, false, true><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, nv, tt, o.inner);
            } else {
                if (any) BasicKernel<true, false, false, true><<<grid, 128, 0, st>>>(e->sc, inO, n, o.hits, o.occ, nv, tt, o.inner);
                else BasicKernel<false, false, false, true><<<grid, 128, 0, st>>>(e->sc, inO, n, o.hits, o.occ, nv, tt, o.inner);
            }
        } else if (e->hasSpheres) {
            if (counters) {
                if (any) BasicKernel<true, true, true><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, o.nodesVisited, o.triTests);
                else BasicKernel<false, true, true><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, o.nodesVisited, o.triTests);
            } else {
                if (any) BasicKernel<true, false, true><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, nullptr, nullptr);
                else BasicKernel<false, false, true><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, nullptr, nullptr);
            }
        } else if (counters) {
            if (any) BasicKernel<true, true><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, o.nodesVisited, o.triTests);
            else BasicKernel<false, true><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, o.nodesVisited, o.triTests);
        } else {
            if (any) BasicKernel<true, false><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, nullptr, nullptr);
            else BasicKernel<false, false><<<grid, 128, 0, st>>>(e->sc, in, n, o.hits, o.occ, nullptr, nullptr);
        }
    } else {
        int minB = (e->wideMinB >= 1 && e->wideMinB <= 3) ? e->wideMinB : 4;
        uint32_t budgetNodes = (uint32_t)((e->maxSmemOptin - 2048) / minB / sizeof(WideNode));
        uint32_t want = e->smemTopNodes >= 0 ? (uint32_t)e->smemTopNodes : 0u;  // measured: L1 beats explicit staging
        uint32_t smemNodes = std::min(std::min(want, budgetNodes), e->sc.topNodes);
        unsigned long long *cursor = scr.cursor;
        CU(cudaMemsetAsync(cursor, 0, kCursorWords * sizeof(unsigned long long), st));
        // ---- binning pre-pass: a ray order that keeps warps coherent (binning.cuh) ----
        // binMode 1: always sort; 2: the device decides per batch (the kernels read the flag); 0: off.
        const uint32_t *perm = nullptr, *permOn = nullptr;
        const bool stepKernel = e->kernel != NNBVH_KERNEL_WIDE || e->hasSpheres || instStep || alphaStep;  // WideKernel: triangles only
        {
            // (an instanced scene's top level is small, its objects are not: no node-count floor there)
            int rc = OrderRays(e, st, in, n, scr, cursor, stepKernel, instStep ? 0u : e->binMinNodes, &perm, &permOn);
            if (rc) return rc;
        }
        if (stepKernel) {
            // ---- StepKernel: fast-slab launch, then the exact launch over the deferred rays ----
            if (n > 0xffffffffull)
                return SetError(NNBVH_ERR_INVALID, "batch too large for one launch");
            StepParams sp{};
            sp.sc = e->sc;
            sp.in = in;
            sp.n = n;
            sp.hits = o.hits;
            sp.occluded = o.occ;
            sp.inner = (instStep && !any) ? o.inner : nullptr;
            sp.perm = perm;
            sp.permOn = permOn;
            StepKnobs(e, &sp.refillBelow, &sp.leafQuorum);
            sp.sphereQuorum = (uint32_t)std::min(32, std::max(1, e->sphereQuorum));
            if (e->stepStats) {
                if (!e->dStats) {
                    CU(cudaMalloc(&e->dStats, 8 * sizeof(unsigned long long)));
                    CU(cudaMemsetAsync(e->dStats, 0, 8 * sizeof(unsigned long long), st));
                }
                sp.stats = e->dStats;
            }
            uint32_t *nDef = reinterpret_cast<uint32_t *>(cursor + 2);
            const bool timed = st == e->stream;  // the device-pointer path; chunk streams are not timed
            if (timed) CU(cudaEventRecord(e->evC, st));
            // any hit: the kernel only writes the occluded rays' 1s (no store left in its node loop)
            if (any) CU(cudaMemsetAsync(o.occ, 0, n, st));
            if (e->boxesRegular && !e->exactOnly) {
                int rc = scr.deferred.Reserve(n);
                if (rc) return rc;
                sp.cursor = cursor;
                sp.deferred = scr.deferred.p;
                sp.nDeferred = nDef;
                if (any) LaunchStep<true, false>(e, st, sp);
                else LaunchStep<false, false>(e, st, sp);
                ++e->launches;
                sp.perm = scr.deferred.p;
                sp.permOn = nullptr;
                sp.nPtr = nDef;
                sp.deferred = nullptr;
                sp.nDeferred = nullptr;
            }
            sp.cursor = cursor + 1;
            if (any) LaunchStep<true, true>(e, st, sp);
            else LaunchStep<false, true>(e, st, sp);
            ++e->launches;
            CU(cudaGetLastError());
            if (timed) {
                CU(cudaEventRecord(e->evD, st));
                e->haveTraversalEvents = true;
            }
            return NNBVH_OK;
        }
        WideParams p;
        p.perm = perm;
        p.sc = e->sc;
        p.in = in;
        p.n = n;
        p.hits = o.hits;
        p.occluded = o.occ;
        p.cursor = cursor;
        p.smemNodes = smemNodes;
        p.refillBelow = (uint32_t)std::min(32, std::max(1, e->refillBelow > 0 ? e->refillBelow : 20));
        p.leafQuorum = (uint32_t)std::min(32, std::max(1, e->leafQuorum > 0 ? e->leafQuorum : 12));
        int rc = any ? LaunchWide<true>(e, st, p, smemNodes * (uint32_t)sizeof(WideNode))
                     : LaunchWide<false>(e, st, p, smemNodes * (uint32_t)sizeof(WideNode));
        if (rc) return rc;
    }
    ++e->launches;
    CU(cudaGetLastError());
    return NNBVH_OK;
}

struct Batch {
    bool any = false, counters = false, soa = false;
    const nnbvh_ray *rays = nullptr;
    const float *planes[7] = {};  // ox oy oz dx dy dz tMax
    uint64_t n = 0;
    Outputs out;
};

static int RunBatch(nnbvh_engine *e, const Batch &b) {
    if (!e) return SetError(NNBVH_ERR_INVALID, "engine is NULL");
    if (!e->haveScene) return SetError(NNBVH_ERR_NOSCENE, "no scene uploaded");
    if (b.n == 0) return NNBVH_OK;
    if (b.soa) {
        for (int k = 0; k < 6; ++k)
            if (!b.planes[k]) return SetError(NNBVH_ERR_INVALID, "SoA ray plane %d is NULL", k);
    } else if (!b.rays) return SetError(NNBVH_ERR_INVALID, "rays is NULL");
    if (b.any ? !b.out.occ : !b.out.hits)
        return SetError(NNBVH_ERR_INVALID, "output buffer is NULL");
    if (b.counters && (!b.out.nodesVisited || !b.out.triTests))
        return SetError(NNBVH_ERR_INVALID, "counter buffer is NULL");
    CU(cudaSetDevice(e->device));

    const bool raysDev = IsDevicePtr(b.soa ? (const void *)b.planes[0] : (const void *)b.rays);
    const bool outDev = IsDevicePtr(b.any ? (const void *)b.out.occ : (const void *)b.out.hits);
    const bool cntDev = !b.counters || IsDevicePtr(b.out.nodesVisited);

    if (raysDev && outDev && cntDev) {
        // ---- everything resident: one launch, asynchronous on the engine's stream ----
        RayInput in{};
        if (b.soa) {
            in.ox = b.planes[0]; in.oy = b.planes[1]; in.oz = b.planes[2];
            in.dx = b.planes[3]; in.dy = b.planes[4]; in.dz = b.planes[5];
            in.tMax = b.planes[6];
        } else in.aos = reinterpret_cast<const float4 *>(b.rays);
        CU(cudaEventRecord(e->evA, e->stream));
        int rc = LaunchTraversal(e, e->stream, b.any, b.counters, in, b.n, b.out, e->scratch);
        if (rc) return rc;
        CU(cudaEventRecord(e->evB, e->stream));
        e->lastMs = -1.f;  // resolved lazily by nnbvh_last_kernel_ms
        return NNBVH_OK;
    }

    // ---- host buffers: chunked H2D -> kernel -> D2H over kSlots streams ----
    const uint64_t chunk = std::max<uint64_t>(b.soa ? e->chunkRaysSoa : e->chunkRays, 1024);
    float total = 0.f;
    // a slot's pooled event pair holds the previous chunk's kernel time until the slot is taken again
    auto collect = [&](Slot &s) {
        if (!s.timed) return;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, s.k0, s.k1) == cudaSuccess) total += ms;
        s.timed = false;
    };
    // one chunk: everything asynchronous on the slot's stream; any error leaves the loop
    auto issue = [&](Slot &s, uint64_t off, uint64_t m) -> int {
        CU(cudaEventSynchronize(s.done));  // slot buffers (and its event pair) free again
        collect(s);
        RayInput in{};
        if (b.soa) {
            if (raysDev) {
                in.ox = b.planes[0] + off; in.oy = b.planes[1] + off; in.oz = b.planes[2] + off;
                in.dx = b.planes[3] + off; in.dy = b.planes[4] + off; in.dz = b.planes[5] + off;
                in.tMax = b.planes[6] ? b.planes[6] + off : nullptr;
            } else {
                int rc = s.soa.Reserve(7 * chunk);
                if (rc) return rc;
                const float **dst[7] = {&in.ox, &in.oy, &in.oz, &in.dx, &in.dy, &in.dz, &in.tMax};
                for (int k = 0; k < 7; ++k) {
                    if (!b.planes[k]) { *dst[k] = nullptr; continue; }
                    CU(cudaMemcpyAsync(s.soa.p + k * chunk, b.planes[k] + off, sizeof(float) * m,
                                       cudaMemcpyHostToDevice, s.stream));
                    *dst[k] = s.soa.p + k * chunk;
                }
            }
        } else if (raysDev) {
            in.aos = reinterpret_cast<const float4 *>(b.rays + off);
        } else {
            int rc = s.rays.Reserve(chunk);
            if (rc) return rc;
            CU(cudaMemcpyAsync(s.rays.p, b.rays + off, sizeof(nnbvh_ray) * m,
                               cudaMemcpyHostToDevice, s.stream));
            in.aos = reinterpret_cast<const float4 *>(s.rays.p);
        }
        Outputs o;
        if (b.any || b.out.occ) {
            if (outDev && b.any) o.occ = b.out.occ + off;
            else if (b.any) { int rc = s.occ.Reserve(chunk); if (rc) return rc; o.occ = s.occ.p; }
        }
        if (!b.any) {
            if (outDev) o.hits = b.out.hits + off;
            else { int rc = s.hits.Reserve(chunk); if (rc) return rc; o.hits = s.hits.p; }
        }
        if (b.out.inner && !b.any) {
            if (outDev) o.inner = b.out.inner + off;
            else { int rc = s.inner.Reserve(chunk); if (rc) return rc; o.inner = s.inner.p; }
        }
        if (b.counters) {
            if (cntDev) { o.nodesVisited = b.out.nodesVisited + off; o.triTests = b.out.triTests + off; }
            else {
                int rc = s.cnt.Reserve(2 * chunk);
                if (rc) return rc;
                o.nodesVisited = s.cnt.p;
                o.triTests = s.cnt.p + chunk;
            }
        }
        CU(cudaEventRecord(s.k0, s.stream));
        int rc = LaunchTraversal(e, s.stream, b.any, b.counters, in, m, o, s.scratch);
        if (rc) return rc;
        CU(cudaEventRecord(s.k1, s.stream));
        s.timed = true;
        if (!outDev) {
            if (b.any) CU(cudaMemcpyAsync(b.out.occ + off, o.occ, m, cudaMemcpyDeviceToHost, s.stream));
            else CU(cudaMemcpyAsync(b.out.hits + off, o.hits, sizeof(nnbvh_hit) * m,
                                    cudaMemcpyDeviceToHost, s.stream));
        }
        if (!outDev && !b.any && b.out.inner)
            CU(cudaMemcpyAsync(b.out.inner + off, o.inner, sizeof(uint32_t) * m, cudaMemcpyDeviceToHost, s.stream));
        if (b.counters && !cntDev) {
            CU(cudaMemcpyAsync(b.out.nodesVisited + off, o.nodesVisited, sizeof(uint32_t) * m,
                               cudaMemcpyDeviceToHost, s.stream));
            CU(cudaMemcpyAsync(b.out.triTests + off, o.triTests, sizeof(uint32_t) * m,
                               cudaMemcpyDeviceToHost, s.stream));
        }
        CU(cudaEventRecord(s.done, s.stream));
        return NNBVH_OK;
    };
    // Chunk schedule.  The first chunk's upload and the last chunk's kernel + download are the part of a call
    // that nothing overlaps, so the schedule ramps up (chunk/4, chunk/2, chunk, ...) and down again at the end
    // (NNBVH_CHUNK_RAMP=0: equal chunks).
    int rcOut = NNBVH_OK;
    int idx = 0;
    const uint64_t small = std::max<uint64_t>(chunk / 4, 1024);
    for (uint64_t off = 0; off < b.n && rcOut == NNBVH_OK; ++idx) {
        uint64_t m = chunk;
        if (e->chunkRamp) {
            if (idx == 0) m = small;
            else if (idx == 1) m = std::max<uint64_t>(chunk / 2, 1024);
            const uint64_t left = b.n - off;
            if (left < 2 * chunk) m = std::min(m, std::max(small, (left + 1) / 2));  // ... halving towards the end
        }
        m = std::min(m, b.n - off);
        rcOut = issue(e->slots[idx % e->nSlots], off, m);
        off += m;
    }
    // Also on an error: nothing may still be copying into the caller's buffers when this returns.
    for (Slot &s : e->slots) {
        cudaError_t ce = cudaStreamSynchronize(s.stream);
        if (ce != cudaSuccess && rcOut == NNBVH_OK)
            rcOut = SetError(NNBVH_ERR_CUDA, "cudaStreamSynchronize failed: %s", cudaGetErrorString(ce));
        collect(s);
    }
    cudaGetLastError();
    e->lastMs = total;
    return rcOut;
}

}  // namespace nnbvh

// ---------------------------------------------------------------------------
// Public entry points
// ---------------------------------------------------------------------------
extern "C" int nnbvh_intersect_closest(nnbvh_engine *e, const nnbvh_ray *rays, uint64_t n,
                                       nnbvh_hit *hits) {
    Batch b;
    b.rays = rays;
    b.n = n;
    b.out.hits = hits;
    return RunBatch(e, b);
}

extern "C" int nnbvh_intersect_closest_instanced(nnbvh_engine *e, const nnbvh_ray *rays, uint64_t n,
                                                 nnbvh_hit *hits, uint32_t *innerPrim) {
    if (e && !e->hasInstances && !(e->general && e->sc.instances))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_instanced: the scene has no instances");
    Batch b;
    b.rays = rays;
    b.n = n;
    b.out.hits = hits;
    b.out.inner = innerPrim;
    if (innerPrim && IsDevicePtr(innerPrim) != IsDevicePtr(hits))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_instanced: hits and innerPrim must live on the same side");
    return RunBatch(e, b);
}

extern "C" int nnbvh_intersect_any(nnbvh_engine *e, const nnbvh_ray *rays, uint64_t n,
                                   uint8_t *occluded) {
    Batch b;
    b.any = true;
    b.rays = rays;
    b.n = n;
    b.out.occ = occluded;
    return RunBatch(e, b);
}

static void FillPlanes(Batch &b, const float *ox, const float *oy, const float *oz,
                       const float *dx, const float *dy, const float *dz, const float *tMax) {
    b.soa = true;
    b.planes[0] = ox; b.planes[1] = oy; b.planes[2] = oz;
    b.planes[3] = dx; b.planes[4] = dy; b.planes[5] = dz;
    b.planes[6] = tMax;
}

extern "C" int nnbvh_intersect_closest_soa(nnbvh_engine *e, const float *ox, const float *oy,
                                           const float *oz, const float *dx, const float *dy,
                                           const float *dz, const float *tMax, uint64_t n,
                                           nnbvh_hit *hits) {
    Batch b;
    FillPlanes(b, ox, oy, oz, dx, dy, dz, tMax);
    b.n = n;
    b.out.hits = hits;
    return RunBatch(e, b);
}

extern "C" int nnbvh_intersect_closest_instanced_soa(nnbvh_engine *e, const float *ox, const float *oy,
                                                     const float *oz, const float *dx, const float *dy,
                                                     const float *dz, const float *tMax, uint64_t n,
                                                     nnbvh_hit *hits, uint32_t *innerPrim) {
    if (e && !e->hasInstances && !(e->general && e->sc.instances))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_instanced_soa: the scene has no instances");
    if (innerPrim && IsDevicePtr(innerPrim) != IsDevicePtr(hits))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_instanced_soa: hits and innerPrim must live on the same side");
    Batch b;
    FillPlanes(b, ox, oy, oz, dx, dy, dz, tMax);
    b.n = n;
    b.out.hits = hits;
    b.out.inner = innerPrim;
    return RunBatch(e, b);
}

extern "C" int nnbvh_intersect_any_soa(nnbvh_engine *e, const float *ox, const float *oy,
                                       const float *oz, const float *dx, const float *dy,
                                       const float *dz, const float *tMax, uint64_t n,
                                       uint8_t *occluded) {
    Batch b;
    FillPlanes(b, ox, oy, oz, dx, dy, dz, tMax);
    b.any = true;
    b.n = n;
    b.out.occ = occluded;
    return RunBatch(e, b);
}

extern "C" int nnbvh_intersect_counters(nnbvh_engine *e, const nnbvh_ray *rays, uint64_t n,
                                        int32_t anyHit, nnbvh_hit *hits, uint8_t *occluded,
                                        uint32_t *nodesVisited, uint32_t *triTests) {
    Batch b;
    b.any = anyHit != 0;
    b.counters = true;
    b.rays = rays;
    b.n = n;
    b.out.hits = hits;
    b.out.occ = occluded;
    b.out.nodesVisited = nodesVisited;
    b.out.triTests = triTests;
    return RunBatch(e, b);
}

// ---------------------------------------------------------------------------
// Known-answer hooks: the device versions of IntersectTriangle / Bounds3f::IntersectP
// ---------------------------------------------------------------------------
namespace nnbvh {

__global__ void TestTrianglesKernel(const float4 *__restrict__ rays, const float *__restrict__ tris9,
                                    uint64_t n, uint8_t *__restrict__ hit, float *__restrict__ out4) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = rays[2 * i], b = rays[2 * i + 1];
    RayConst rc = MakeRayConst(a.x, a.y, a.z, b.x, b.y, b.z);
    const float *p = tris9 + 9 * i;
    float t = 0, b0 = 0, b1 = 0, b2 = 0;
    bool h = false;
    if (!TriangleDegenerate(p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8]))
        h = TriangleTest(rc, a.w, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8], &t, &b0,
                         &b1, &b2);
    hit[i] = h ? 1 : 0;
    out4[4 * i + 0] = h ? t : 0.f;
    out4[4 * i + 1] = h ? b0 : 0.f;
    out4[4 * i + 2] = h ? b1 : 0.f;
    out4[4 * i + 3] = h ? b2 : 0.f;
}

__global__ void TestBoxesKernel(const float4 *__restrict__ rays, const float *__restrict__ boxes6,
                                uint64_t n, uint8_t *__restrict__ hit) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = rays[2 * i], b = rays[2 * i + 1];
    RayConst rc = MakeRayConst(a.x, a.y, a.z, b.x, b.y, b.z);
    float box[6];
    for (int k = 0; k < 6; ++k) box[k] = boxes6[6 * i + k];
    hit[i] = SlabTest(rc, box, a.w) ? 1 : 0;
}

}  // namespace nnbvh

extern "C" int nnbvh_test_triangles(nnbvh_engine *e, const nnbvh_ray *rays, const float *tris9,
                                    uint64_t n, uint8_t *hit, float *out4) {
    if (!e || !rays || !tris9 || !hit || !out4)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_test_triangles: NULL argument");
    if (n == 0) return NNBVH_OK;
    CU(cudaSetDevice(e->device));
    cudaStream_t st = e->stream;
    DevTmp<float4> dR;
    DevTmp<float> dT, dO;
    DevTmp<uint8_t> dH;
    CU(StageIn(dR, rays, 2 * n, st));
    CU(StageIn(dT, tris9, 9 * n, st));
    CU(dO.Alloc(4 * n));
    CU(dH.Alloc(n));
    TestTrianglesKernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(dR.p, dT.p, n, dH.p, dO.p);
    ++e->launches;
    cudaError_t ce = cudaGetLastError();
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(hit, dH.p, n, cudaMemcpyDeviceToHost, st);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(out4, dO.p, n * 16, cudaMemcpyDeviceToHost, st);
    cudaError_t se = cudaStreamSynchronize(st);  // before the temporaries are freed, error or not
    CU(ce);
    CU(se);
    return NNBVH_OK;
}

extern "C" int nnbvh_test_boxes(nnbvh_engine *e, const nnbvh_ray *rays, const float *boxes6,
                                uint64_t n, uint8_t *hit) {
    if (!e || !rays || !boxes6 || !hit)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_test_boxes: NULL argument");
    if (n == 0) return NNBVH_OK;
    CU(cudaSetDevice(e->device));
    cudaStream_t st = e->stream;
    DevTmp<float4> dR;
    DevTmp<float> dB;
    DevTmp<uint8_t> dH;
    CU(StageIn(dR, rays, 2 * n, st));
    CU(StageIn(dB, boxes6, 6 * n, st));
    CU(dH.Alloc(n));
    TestBoxesKernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(dR.p, dB.p, n, dH.p);
    ++e->launches;
    cudaError_t ce = cudaGetLastError();
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(hit, dH.p, n, cudaMemcpyDeviceToHost, st);
    cudaError_t se = cudaStreamSynchronize(st);
    CU(ce);
    CU(se);
    return NNBVH_OK;
}

namespace nnbvh {
__global__ void TestSpheresKernel(const float4 *__restrict__ rays, const SphereDev *__restrict__ spheres,
                                  uint64_t n, uint8_t *__restrict__ hit, float4 *__restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = rays[2 * i], b = rays[2 * i + 1];
    SphereHit sh{0.f, 0.f, 0.f, 0.f};
    const bool ok = SphereTest(spheres + i, a.x, a.y, a.z, b.x, b.y, b.z, a.w, &sh);
    hit[i] = ok ? 1 : 0;
    out[i] = ok ? make_float4(sh.t, sh.px, sh.py, sh.pz) : make_float4(0.f, 0.f, 0.f, 0.f);
}
__global__ void TestAtan2Kernel(const float *__restrict__ y, const float *__restrict__ x, uint64_t n,
                                float *__restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = Atan2fHost(y[i], x[i]);
}
}  // namespace nnbvh

extern "C" int nnbvh_test_atan2(nnbvh_engine *e, const float *y, const float *x, uint64_t n, float *out) {
    if (!e || !y || !x || !out) return SetError(NNBVH_ERR_INVALID, "nnbvh_test_atan2: NULL argument");
    if (n == 0) return NNBVH_OK;
    CU(cudaSetDevice(e->device));
    cudaStream_t st = e->stream;
    DevTmp<float> d;
    CU(d.Alloc(3 * n));
    cudaError_t ce = cudaMemcpyAsync(d.p, y, n * 4, cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(d.p + n, x, n * 4, cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) {
        TestAtan2Kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d.p, d.p + n, n, d.p + 2 * n);
        ++e->launches;
        ce = cudaGetLastError();
    }
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(out, d.p + 2 * n, n * 4, cudaMemcpyDeviceToHost, st);
    cudaError_t se = cudaStreamSynchronize(st);
    CU(ce);
    CU(se);
    return NNBVH_OK;
}

extern "C" int nnbvh_test_spheres(nnbvh_engine *e, const nnbvh_ray *rays, const nnbvh_sphere *spheres,
                                  uint64_t n, uint8_t *hit, float *out4) {
    if (!e || !rays || !spheres || !hit || !out4)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_test_spheres: NULL argument");
    if (n == 0) return NNBVH_OK;
    CU(cudaSetDevice(e->device));
    std::vector<SphereDev> sd(n);
    for (uint64_t i = 0; i < n; ++i) {
        int rc = MakeSphereDev(spheres[i], (uint32_t)i, &sd[i]);
        if (rc) return rc;
    }
    cudaStream_t st = e->stream;
    DevTmp<float4> dR, dO;
    DevTmp<SphereDev> dS;
    DevTmp<uint8_t> dH;
    CU(StageIn(dR, rays, 2 * n, st));
    cudaError_t ce = StageIn(dS, sd.data(), n, st);
    if (ce == cudaSuccess) ce = dH.Alloc(n);
    if (ce == cudaSuccess) ce = dO.Alloc(n);
    if (ce == cudaSuccess) {
        TestSpheresKernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(dR.p, dS.p, n, dH.p, dO.p);
        ++e->launches;
        ce = cudaGetLastError();
    }
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(hit, dH.p, n, cudaMemcpyDeviceToHost, st);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(out4, dO.p, n * 16, cudaMemcpyDeviceToHost, st);
    cudaError_t se = cudaStreamSynchronize(st);  // `sd` and the temporaries outlive every copy
    CU(ce);
    CU(se);
    return NNBVH_OK;
}

// ---------------------------------------------------------------------------
// HLBVH with the data-parallel stages on the device (hlbvh.cuh)
// ---------------------------------------------------------------------------
namespace {
struct HlbvhBufs {
    float *tris = nullptr, *box = nullptr;
    uint32_t *u32 = nullptr;  // one arena, see the layout in nnbvh_build_hlbvh_device
    nnbvh_node *nodes = nullptr;
    ~HlbvhBufs() {
        cudaFree(tris); cudaFree(box); cudaFree(u32); cudaFree(nodes);
    }
};
}  // namespace

extern "C" int nnbvh_build_hlbvh_device(nnbvh_engine *e, const float *triVerts9, uint32_t nTris,
                                        uint32_t maxPrimsInNode, nnbvh_node *nodesOut,
                                        uint32_t *nNodesOut, uint32_t *orderOut) {
    if (!e) return SetError(NNBVH_ERR_INVALID, "nnbvh_build_hlbvh_device: engine is NULL");
    if (!triVerts9 || !nodesOut || !nNodesOut || !orderOut)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_hlbvh_device: NULL argument");
    if (nTris == 0) return SetError(NNBVH_ERR_INVALID, "nnbvh_build_hlbvh_device: no primitives");
    if (nTris > (1u << 30))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_hlbvh_device: too many primitives");
    const uint32_t maxLeaf = std::min<uint32_t>(255u, maxPrimsInNode);  // aggregates.cpp:149
    CU(cudaSetDevice(e->device));
    cudaStream_t st = e->stream;
    const uint32_t n = nTris;
    const uint32_t nBlk = (n + 255) / 256, nSortBlk = (n + kSortTile - 1) / kSortTile;
    const size_t nHist = 256 * (size_t)nSortBlk;
    const size_t nTiles = (std::max<size_t>(nHist, (size_t)n + 1) + kScanTile - 1) / kScanTile;
    // layout of the u32 arena
    size_t at = 0;
    auto take = [&](size_t k) { size_t r = at; at += (k + 3) & ~(size_t)3; return r; };
    const size_t oCb = take(8), oHead = take(kMaxTreelets), oStarts = take(kMaxTreelets),
                 oOffs = take(kMaxTreelets + 1), oCount = take(1), oOv = take(1), oCodeA = take(n),
                 oCodeB = take(n), oPrimA = take(n), oPrimB = take(n), oHist = take(nHist),
                 oTile = take(nTiles), oMask = take(n), oLeafN = take(n), oRangeL = take(n),
                 oS = take((size_t)n + 1), oParent = take(2 * (size_t)n), oVisit = take(2 * (size_t)n);
    HlbvhBufs b;
    CU(cudaMalloc(&b.tris, 36 * (size_t)n));
    CU(cudaMalloc(&b.box, 24 * (size_t)n));
    CU(cudaMalloc(&b.u32, 4 * at));
    CU(cudaMalloc(&b.nodes, sizeof(nnbvh_node) * 2 * (size_t)n));  // a tree over n leaves: < 2n nodes
    uint32_t *u = b.u32;
    CU(cudaMemcpyAsync(b.tris, triVerts9, 36 * (size_t)n, cudaMemcpyHostToDevice, st));
    CU(cudaEventRecord(e->evA, st));
    CU(cudaMemsetAsync(u + oCb, 0xff, 12, st));
    CU(cudaMemsetAsync(u + oCb + 3, 0, 12, st));
    CU(cudaMemsetAsync(u + oHead, 0xff, 4 * kMaxTreelets, st));
    CU(cudaMemsetAsync(u + oOv, 0, 4, st));
    CU(cudaMemsetAsync(u + oMask, 0, 4 * (size_t)n, st));
    CU(cudaMemsetAsync(u + oVisit, 0, 8 * (size_t)n, st));
    PrimBoundsKernel<<<nBlk, 256, 0, st>>>(b.tris, n, b.box, u + oCb);
    MortonKernel<<<nBlk, 256, 0, st>>>(b.box, n, u + oCb, u + oCodeA, u + oPrimA);
    e->launches += 2;
    uint32_t *cIn = u + oCodeA, *cOut = u + oCodeB, *pIn = u + oPrimA, *pOut = u + oPrimB;
    const unsigned nHistTiles = (unsigned)((nHist + kScanTile - 1) / kScanTile);
    for (int shift = 0; shift < 30; shift += 8) {
        RadixCountKernel<8><<<nSortBlk, 256, 0, st>>>(cIn, n, shift, u + oHist);
        BinTileSumKernel<<<nHistTiles, 1024, 0, st>>>(u + oHist, (uint32_t)nHist, u + oTile);
        BinTileScanKernel<<<nHistTiles, 1024, 0, st>>>(u + oHist, (uint32_t)nHist, u + oTile);
        RadixScatterKernel<8><<<nSortBlk, 256, 0, st>>>(cIn, pIn, n, shift, u + oHist, cOut, pOut);
        e->launches += 4;
        std::swap(cIn, cOut);
        std::swap(pIn, pOut);
    }
    TreeletHeadKernel<<<nBlk, 256, 0, st>>>(cIn, n, u + oHead);
    TreeletListKernel<<<1, 32, 0, st>>>(u + oHead, u + oStarts, u + oCount);
    LbvhTopologyKernel<<<nBlk, 256, 0, st>>>(cIn, n, maxLeaf, u + oMask, u + oLeafN, u + oRangeL,
                                             reinterpret_cast<int *>(u + oOv));
    const unsigned nSTiles = (unsigned)(((size_t)n + 1 + kScanTile - 1) / kScanTile);
    LbvhCountKernel<<<(n + 1 + 255) / 256, 256, 0, st>>>(u + oMask, n, u + oS);
    BinTileSumKernel<<<nSTiles, 1024, 0, st>>>(u + oS, n + 1, u + oTile);
    BinTileScanKernel<<<nSTiles, 1024, 0, st>>>(u + oS, n + 1, u + oTile);
    LbvhNodesKernel<<<nBlk, 256, 0, st>>>(cIn, n, u + oMask, u + oLeafN, u + oRangeL, u + oS,
                                          u + oHead, b.nodes, u + oParent);
    LbvhBoundsKernel<<<nBlk, 256, 0, st>>>(cIn, pIn, b.box, n, u + oMask, u + oLeafN, u + oS,
                                           u + oHead, u + oParent, u + oVisit, b.nodes);
    TreeletOffsetsKernel<<<(kMaxTreelets + 1 + 255) / 256, 256, 0, st>>>(u + oStarts, u + oCount,
                                                                        u + oS, n, u + oOffs);
    e->launches += 9;
    CU(cudaGetLastError());
    CU(cudaEventRecord(e->evB, st));
    // control block back: count, overflow, then starts and offs
    uint32_t hCount = 0;
    int hOv = 0;
    CU(cudaMemcpyAsync(&hCount, u + oCount, 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&hOv, u + oOv, 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaEventElapsedTime(&e->lastMs, e->evA, e->evB));
    if (hOv)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_hlbvh_device: a leaf holds >= 65536 "
                                           "primitives (the reference CHECKs nPrimitives < 65536)");
    std::vector<uint32_t> starts(hCount), offs(hCount + 1);
    CU(cudaMemcpyAsync(starts.data(), u + oStarts, 4 * (size_t)hCount, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(offs.data(), u + oOffs, 4 * (size_t)(hCount + 1), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const uint32_t total = offs[hCount];
    std::vector<nnbvh_node> dense(total);
    CU(cudaMemcpyAsync(dense.data(), b.nodes, sizeof(nnbvh_node) * (size_t)total,
                       cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(orderOut, pIn, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    std::vector<TreeletRef> tr(hCount);
    for (uint32_t k = 0; k < hCount; ++k)
        tr[k] = TreeletRef{starts[k], (k + 1 < hCount ? starts[k + 1] : n) - starts[k], offs[k],
                           offs[k + 1] - offs[k]};
    return AssembleHLBVH(dense.data(), tr.data(), hCount, nodesOut, nNodesOut);
}

// ---------------------------------------------------------------------------
// Result window on a peer GPU (the exchange step of the sharded path): the caller's
// rank allocates the global result array, the other ranks map it through CUDA IPC and deliver
// their slice with a copy engine over NVLink -- no SM is taken from the traversal kernels, which
// fill the GPU with persistent CTAs and would otherwise serialise with a collective's kernel.
// ---------------------------------------------------------------------------
extern "C" int nnbvh_peer_alloc(uint64_t bytes, void **devPtr, void *handle64) {
    if (!devPtr || !handle64) return SetError(NNBVH_ERR_INVALID, "nnbvh_peer_alloc: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    void *p = nullptr;
    CU(cudaMalloc(&p, std::max<uint64_t>(bytes, 1)));
    cudaIpcMemHandle_t h;
    cudaError_t ce = cudaIpcGetMemHandle(&h, p);
    if (ce != cudaSuccess) {
        cudaFree(p);
        return SetError(NNBVH_ERR_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(ce));
    }
    std::memcpy(handle64, &h, 64);
    *devPtr = p;
    return NNBVH_OK;
}

extern "C" int nnbvh_peer_open(const void *handle64, void **devPtr) {
    if (!devPtr || !handle64) return SetError(NNBVH_ERR_INVALID, "nnbvh_peer_open: NULL argument");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    CU(cudaIpcOpenMemHandle(devPtr, h, cudaIpcMemLazyEnablePeerAccess));
    return NNBVH_OK;
}

extern "C" int nnbvh_peer_close(void *devPtr) {
    if (devPtr) CU(cudaIpcCloseMemHandle(devPtr));
    return NNBVH_OK;
}

extern "C" int nnbvh_peer_free(void *devPtr) {
    if (devPtr) CU(cudaFree(devPtr));
    return NNBVH_OK;
}

extern "C" int nnbvh_peer_copy(void *dst, const void *src, uint64_t bytes, void *cudaStream) {
    if (bytes == 0) return NNBVH_OK;
    if (!dst || !src) return SetError(NNBVH_ERR_INVALID, "nnbvh_peer_copy: NULL argument");
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, static_cast<cudaStream_t>(cudaStream)));
    return NNBVH_OK;
}

extern "C" int nnbvh_pointer_kind(const void *p, int *kind) {
    if (!kind) return SetError(NNBVH_ERR_INVALID, "nnbvh_pointer_kind: kind is NULL");
    *kind = 0;
    if (!p) return NNBVH_OK;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();  // no device, or a pointer CUDA has never seen: host memory
        return NNBVH_OK;
    }
    *kind = a.type == cudaMemoryTypeDevice ? 1 : a.type == cudaMemoryTypeManaged ? 2 : 0;
    return NNBVH_OK;
}

extern "C" int nnbvh_copy_to_host(void *dstHost, const void *src, uint64_t bytes) {
    if (bytes == 0) return NNBVH_OK;
    if (!dstHost || !src) return SetError(NNBVH_ERR_INVALID, "nnbvh_copy_to_host: NULL argument");
    CU(cudaMemcpy(dstHost, src, bytes, cudaMemcpyDefault));
    return NNBVH_OK;
}

// ---------------------------------------------------------------------------
// Read-bandwidth probe: the denominator SURVEY 8(d) asks for next to the HBM figure when the scene
// is L2-resident.  A persistent grid streams a `bytes`-sized buffer `iters` times with the same
// 256-bit read-only loads the traversal kernel uses; small buffers measure L2, large ones HBM.
// ---------------------------------------------------------------------------
namespace nnbvh {
__global__ void __launch_bounds__(256)
ReadProbeKernel(const unsigned char *__restrict__ buf, uint64_t nChunks, uint32_t iters,
                float *__restrict__ sink) {
    float acc = 0.f;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint32_t it = 0; it < iters; ++it)
        for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nChunks; i += stride) {
            const F8 v = Ldg256(buf + 32 * i);
            acc += v.v[0] + v.v[7];
        }
    if (acc == 123.456f) *sink = acc;  // keeps the loads alive
}
}  // namespace nnbvh

extern "C" int nnbvh_read_bandwidth_probe(nnbvh_engine *e, uint64_t bytes, uint32_t iters, float *gbs) {
    if (!e || !gbs) return SetError(NNBVH_ERR_INVALID, "nnbvh_read_bandwidth_probe: NULL argument");
    if (bytes < 32 || iters == 0) return SetError(NNBVH_ERR_INVALID, "nnbvh_read_bandwidth_probe: empty probe");
    CU(cudaSetDevice(e->device));
    unsigned char *buf = nullptr;
    float *sink = nullptr;
    CU(cudaMalloc(&buf, bytes));
    CU(cudaMalloc(&sink, sizeof(float)));
    CU(cudaMemsetAsync(buf, 0, bytes, e->stream));
    const uint64_t nChunks = bytes / 32;
    ReadProbeKernel<<<e->numSMs * 8, 256, 0, e->stream>>>(buf, nChunks, 1, sink);  // warm-up (fills L2)
    CU(cudaEventRecord(e->evA, e->stream));
    ReadProbeKernel<<<e->numSMs * 8, 256, 0, e->stream>>>(buf, nChunks, iters, sink);
    CU(cudaEventRecord(e->evB, e->stream));
    e->launches += 2;
    CU(cudaGetLastError());
    CU(cudaEventSynchronize(e->evB));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, e->evA, e->evB));
    cudaFree(buf);
    cudaFree(sink);
    *gbs = (float)((double)nChunks * 32.0 * iters / (ms * 1e-3) / 1e9);
    return NNBVH_OK;
}

// ---------------------------------------------------------------------------
// nnbvh_intersect_closest_noalloc: the closest-hit StepKernel path for callers that cannot allocate
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The StepKernel path needs per-launch scratch: the ray cursors / deferred-ray count (kCursorWords
// 64-bit words) and the list of rays deferred to the exact slab test (one uint32 per ray).  The
// ordinary entry points keep them in cudaMalloc'ed buffers that grow on demand (Scratch); this one
// keeps them in ONE file-scope device array of fixed size, so it calls nothing that allocates:
// cudaMemsetAsync of the cursor words, two kernel launches.  Every GPU that loads libnnbvh.so holds
// the array (256 B + 4 B x NNBVH_NOALLOC_MAX_RAYS = 4 MiB + 256 B) until the process exits.
//
// All calls on one device share the array: calls that overlap on the GPU (two streams, two engines,
// two threads) race on the cursors.  Calls ordered on one stream are safe; the cursor words are
// cleared at the start of each call, the deferred list needs no clearing (only entries below the
// freshly counted nDeferred are read).
// ---------------------------------------------------------------------------
namespace nnbvh {
constexpr size_t kNoAllocDeferredAt = 256;  // bytes: cursor words first, the deferred list after them
static_assert(kCursorWords * sizeof(unsigned long long) <= kNoAllocDeferredAt, "cursor words fit the header");
static __device__ __align__(256) unsigned char g_noAllocScratch[kNoAllocDeferredAt +
                                                                sizeof(uint32_t) * (size_t)NNBVH_NOALLOC_MAX_RAYS];
// In host code `g_noAllocScratch` is only a stand-in for the symbol, NOT a device address: the address
// is asked for once per device (cudaGetSymbolAddress, that device current) and kept here.
constexpr int kNoAllocMaxDevices = 64;
static void *g_noAllocPtr[kNoAllocMaxDevices] = {};
}  // namespace nnbvh

static int NoAllocSetup(int device) {
    if (device < 0 || device >= kNoAllocMaxDevices) return NNBVH_OK;  // the entry point reports it
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (g_noAllocPtr[device]) return NNBVH_OK;
    void *ptr = nullptr;
    CU(cudaGetSymbolAddress(&ptr, g_noAllocScratch));
    if (!ptr) return SetError(NNBVH_ERR_CUDA, "cudaGetSymbolAddress(g_noAllocScratch) returned NULL on device %d", device);
    // Load the kernels the entry point launches now (CUDA loads functions lazily on first use, and
    // that load may allocate): default instantiations; an engine created with NNBVH_WIDE_MINB set
    // loads its own on its first ordinary batch.
    cudaFuncAttributes fa;
    CU(cudaFuncGetAttributes(&fa, StepKernel<false, false, 5>));
    CU(cudaFuncGetAttributes(&fa, StepKernel<false, true, 5>));
    CU(cudaFuncGetAttributes(&fa, StepKernel<false, false, 4, false, true>));
    CU(cudaFuncGetAttributes(&fa, StepKernel<false, true, 4, false, true>));
    g_noAllocPtr[device] = ptr;
    return NNBVH_OK;
}

extern "C" int nnbvh_intersect_closest_noalloc(nnbvh_engine *e, const nnbvh_ray *raysDev, uint64_t n,
                                               nnbvh_hit *hitsDev) {
    if (!e) return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_noalloc: engine is NULL");
    if (!e->haveScene) return SetError(NNBVH_ERR_NOSCENE, "no scene uploaded");
    if (n == 0) return NNBVH_OK;
    if (n > (uint64_t)NNBVH_NOALLOC_MAX_RAYS)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_noalloc: %llu rays exceed NNBVH_NOALLOC_MAX_RAYS = %u "
                        "(the static scratch array holds one deferred-list entry per ray); split the batch or "
                        "use nnbvh_intersect_closest", (unsigned long long)n, (unsigned)NNBVH_NOALLOC_MAX_RAYS);
    if (!raysDev || !hitsDev) return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_noalloc: NULL rays / hits");
    if (e->hasAlpha || e->hasInstances || e->general)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_noalloc: scenes with alpha-tested or instanced "
                                           "primitives run other kernels; use nnbvh_intersect_closest");
    if (e->device < 0 || e->device >= kNoAllocMaxDevices || !g_noAllocPtr[e->device])
        return SetError(NNBVH_ERR_CUDA, "nnbvh_intersect_closest_noalloc: no scratch address for device %d", e->device);
    CU(cudaSetDevice(e->device));
    if (!IsDevicePtr(raysDev) || !IsDevicePtr(hitsDev))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_intersect_closest_noalloc: rays and hits must be device pointers "
                                           "(staging host buffers would need an allocation)");
    unsigned char *base = static_cast<unsigned char *>(g_noAllocPtr[e->device]);
    unsigned long long *cursor = reinterpret_cast<unsigned long long *>(base);
    uint32_t *deferred = reinterpret_cast<uint32_t *>(base + kNoAllocDeferredAt);
    cudaStream_t st = e->stream;
    CU(cudaMemsetAsync(cursor, 0, kCursorWords * sizeof(unsigned long long), st));
    StepParams sp{};
    sp.sc = e->sc;
    sp.in.aos = reinterpret_cast<const float4 *>(raysDev);
    sp.n = n;
    sp.hits = hitsDev;
    StepKnobs(e, &sp.refillBelow, &sp.leafQuorum);
    sp.sphereQuorum = (uint32_t)std::min(32, std::max(1, e->sphereQuorum));
    if (e->boxesRegular && !e->exactOnly) {  // fast-slab launch; rays it cannot take go to the deferred list
        sp.cursor = cursor;
        sp.deferred = deferred;
        sp.nDeferred = reinterpret_cast<uint32_t *>(cursor + 2);
        LaunchStep<false, false>(e, st, sp);
        ++e->launches;
        sp.perm = deferred;
        sp.nPtr = sp.nDeferred;
        sp.deferred = nullptr;
        sp.nDeferred = nullptr;
    }
    sp.cursor = cursor + 1;
    LaunchStep<false, true>(e, st, sp);  // exact launch over the deferred rays (or over all of them)
    ++e->launches;
    CU(cudaGetLastError());
    return NNBVH_OK;
}
