/* include/nnbvh.h -- C ABI of libnnbvh.so, the B200-native batch ray-query engine.
 *
 * Drop-in boundary for ONE path of the reference (Wil-Lee/NN_BVH = pbrt-v4 +
 * machine_learning/): closest-hit / any-hit traversal of the flattened BVH plus
 * the watertight ray-triangle test.  The reference has no FFI for this path
 * (SURVEY.md section 8(b)); these entry points are what a binding would wrap.
 * Paths below are relative to /root/reference/src/pbrt.
 *
 * Conventions: every function returns 0 on success or a negative nnbvh_status;
 * nothing aborts or throws across the boundary (the reference LOG_FATALs,
 * util/check.h:36).  nnbvh_last_error() gives the message for the calling
 * thread.  Ray / hit / output pointers may be host (pageable or pinned) or
 * device pointers; device pointers are detected with cudaPointerGetAttributes
 * and used in place.  There is NO CPU fallback: without a CUDA device every
 * compute entry point fails with NNBVH_ERR_CUDA.
 *
 * Threading: an nnbvh_engine is NOT re-entrant.  It owns one launch stream, one set of scratch
 * buffers and one host-pointer pipeline (slots of staging buffers), and it serialises its batches;
 * calls on ONE engine must not overlap in time -- callers with several threads give each thread its
 * own engine (nnbvh_clone shares the uploaded scene) or take a lock (integration/pbrt/gpubvh.h does
 * for the single-ray Primitive surface).  DIFFERENT engines may be used from different threads at
 * once.  nnbvh_last_error() is per thread.  The host builders (nnbvh_build_bvh*) are self-contained
 * and may run concurrently.  nnbvh_intersect_closest_noalloc additionally shares one static scratch
 * array between all engines of a device (see its comment).
 */
#ifndef NNBVH_H
#define NNBVH_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum nnbvh_status {
    NNBVH_OK = 0,
    NNBVH_ERR_INVALID = -1,   /* bad argument / malformed tree */
    NNBVH_ERR_CUDA = -2,      /* CUDA runtime error or no device */
    NNBVH_ERR_NOSCENE = -3,   /* query before nnbvh_upload_scene */
    NNBVH_ERR_DEPTH = -4,     /* traversal stack would exceed 64 entries
                                 (cpu/aggregates.cpp:538 has no such check) */
    NNBVH_ERR_NOMEM = -5
} nnbvh_status;

/* Byte-identical to `struct alignas(32) LinearBVHNode`
 * (cpu/aggregates.cpp:128-137): pre-order, first child at idx+1. */
typedef struct nnbvh_node {
    float bounds[6];       /* pMin.xyz, pMax.xyz */
    int32_t offset;        /* leaf: primitivesOffset; interior: secondChildOffset */
    uint16_t nPrimitives;  /* 0 -> interior */
    uint8_t axis;
    uint8_t pad;
} nnbvh_node;

/* Ray {o, d, time} (ray.h:35-38) + the tMax argument of Intersect/IntersectP. */
typedef struct nnbvh_ray {
    float ox, oy, oz, tMax, dx, dy, dz, time;
} nnbvh_ray;

/* What survives of ShapeIntersection (shapes.h:93-97): ordered primitive index
 * (NNBVH_MISS = no hit), tHit and the TriangleIntersection barycentrics
 * (shapes.h:820-824).  SurfaceInteraction is rebuilt on the host from these
 * (Triangle::InteractionFromIntersection, shapes.h:884). */
typedef struct nnbvh_hit {
    uint32_t prim;
    float t, b0, b1, b2;
} nnbvh_hit;

#define NNBVH_MISS 0xFFFFFFFFu

/* BVHAggregate::SplitMethod (cpu/aggregates.h:31). */
typedef enum nnbvh_split { NNBVH_SPLIT_SAH = 0, NNBVH_SPLIT_HLBVH = 1,
                           NNBVH_SPLIT_MIDDLE = 2, NNBVH_SPLIT_EQUAL_COUNTS = 3 } nnbvh_split;

typedef enum nnbvh_kernel {
    NNBVH_KERNEL_AUTO = 0,   /* engine default (the fast path) */
    NNBVH_KERNEL_BASIC = 1,  /* one thread per ray over the 32-byte nodes, statement
                                order of cpu/aggregates.cpp:529-624 */
    NNBVH_KERNEL_WIDE = 2,   /* persistent warps over 64-byte two-child nodes (first version) */
    NNBVH_KERNEL_STEP = 3    /* persistent warps, one straight-line step per lane per
                                iteration, 256-bit node loads (the default) */
} nnbvh_kernel;

typedef struct nnbvh_options {
    int32_t device;          /* CUDA device ordinal; -1 = current device */
    int32_t kernel;          /* nnbvh_kernel */
    int32_t smem_top_nodes;  /* wide kernel: top-of-tree nodes staged in shared
                                memory; -1 = default, 0 = none */
    int32_t reserved[5];
} nnbvh_options;

typedef struct nnbvh_engine nnbvh_engine;

/* ---- host-side tree construction (no GPU needed) ------------------------- */

/* Replaces BVHAggregate::BVHAggregate + buildRecursive + flattenBVH
 * (cpu/aggregates.cpp:140-387, 505-522) for a triangle soup: same splits, same
 * leaf order, byte-identical nodes (the reference on ONE thread; its parallel
 * build orders leaves nondeterministically, :359-370).
 * triVerts9: nTris*9 floats, original order.  nodesOut: capacity 2*nTris-1.
 * orderOut[i] = original index of ordered primitive i. */
int nnbvh_build_bvh(const float *triVerts9, uint32_t nTris, int32_t maxPrimsInNode,
                    int32_t splitMethod, int32_t nThreads, nnbvh_node *nodesOut,
                    uint32_t *nNodesOut, uint32_t *orderOut);

/* The same builder over arbitrary primitives: bounds6 = nPrims * {pMin xyz, pMax xyz}, i.e. what
 * Primitive::Bounds() returns for each entry of the std::vector<Primitive> handed to
 * BVHAggregate (cpu/aggregates.cpp:150-153).  Triangles: Triangle::Bounds (shapes.cpp:294-301);
 * spheres: nnbvh_sphere_bounds. */
int nnbvh_build_bvh_bounds(const float *bounds6, uint32_t nPrims, int32_t maxPrimsInNode,
                           int32_t splitMethod, int32_t nThreads, nnbvh_node *nodesOut,
                           uint32_t *nNodesOut, uint32_t *orderOut);

/* ---- sphere leaves (Sphere, shapes.h:107-340) ---------------------------------------------
 * What `Shape "sphere"` hands to Sphere::Sphere (shapes.h:117-128): both transforms (rows 0..2 of
 * the 4x4 matrices, row-major; row 3 must be 0 0 0 1) and radius / zmin / zmax / phimax as the
 * scene file gives them (the constructor's clamping is applied by the library).  Spheres with
 * phimax < 360 compare std::atan2's result (shapes.h:196-203): the device follows glibc's
 * single-precision algorithm (csrc/atan2_host.cuh), bit for bit the libm of this image. */
typedef struct nnbvh_sphere {
    float renderFromObject[12];
    float objectFromRender[12];
    float radius, zMin, zMax, phiMax; /* phiMax in degrees */
} nnbvh_sphere;

/* Sphere::Bounds() (shapes.cpp:37-40), on the host: bounds6 = {pMin xyz, pMax xyz}. */
int nnbvh_sphere_bounds(const nnbvh_sphere *s, float *bounds6);

/* Maximum number of simultaneously stacked nodes any ray can need on this tree
 * (the reference's nodesToVisit[64], cpu/aggregates.cpp:538). */
int nnbvh_tree_stack_depth(const nnbvh_node *nodes, uint32_t nNodes, uint32_t *depthOut);

/* ---- engine ---------------------------------------------------------------- */

int nnbvh_create(const nnbvh_options *opts /* may be NULL */, nnbvh_engine **out);
void nnbvh_destroy(nnbvh_engine *e);

/* Stream (cudaStream_t) that all later work of this engine is issued on;
 * NULL = the engine's own stream. */
int nnbvh_set_stream(nnbvh_engine *e, void *cudaStream);

/* nnbvh_build_bvh(..., NNBVH_SPLIT_HLBVH, ...) with the data-parallel stages of
 * BVHAggregate::buildHLBVH (cpu/aggregates.cpp:389-449: primitive and centroid
 * bounds, Morton codes, the radix sort, treelet emission) run on the engine's GPU;
 * buildUpperSAH (:626-723) over the <= 4096 treelet roots stays on the host.
 * Same arguments, byte-identical output.  triVerts9 is a HOST pointer. */
int nnbvh_build_hlbvh_device(nnbvh_engine *e, const float *triVerts9, uint32_t nTris,
                             uint32_t maxPrimsInNode, nnbvh_node *nodesOut,
                             uint32_t *nNodesOut, uint32_t *orderOut);

/* Device time of the traversal launches alone (fast-slab launch + exact launch over the
 * deferred rays) of the last device-pointer batch: nnbvh_last_kernel_ms minus the ray
 * ordering pre-pass.  Blocks until that batch has finished. */
int nnbvh_last_traversal_ms(nnbvh_engine *e, float *ms);

/* ---- sharded runs: result window on the caller's GPU -------------------------
 * One process per GPU.  The rank that wants the results allocates the global array
 * (nnbvh_peer_alloc: device memory + a 64-byte CUDA IPC handle to send to the other
 * processes); every other rank maps it (nnbvh_peer_open) and delivers its slice with
 * nnbvh_peer_copy, a copy-engine transfer over NVLink ordered on `cudaStream` -- no
 * SM is taken from the traversal kernels.  All on the calling thread's current device. */
int nnbvh_peer_alloc(uint64_t bytes, void **devPtr, void *handle64);
int nnbvh_peer_open(const void *handle64, void **devPtr);
int nnbvh_peer_close(void *devPtr);
int nnbvh_peer_free(void *devPtr);
int nnbvh_peer_copy(void *dst, const void *src, uint64_t bytes, void *cudaStream);

/* Where a caller's buffer lives (cudaPointerGetAttributes): *kind = 0 host memory (pageable or
 * pinned), 1 device memory, 2 managed memory.  The wavefront backend (nnbvh_wavefront.hpp) uses
 * it to decide whether the host loop after a batch can read a queue's planes in place
 * (the reference's queues are host memory on --wavefront and managed memory on --gpu,
 * wavefront/integrator.cpp:231-234). */
int nnbvh_pointer_kind(const void *p, int *kind);
/* Blocking copy of `bytes` from device (or managed) memory to host memory. */
int nnbvh_copy_to_host(void *dstHost, const void *src, uint64_t bytes);

/* Measurement: read bandwidth (GB/s) of a `bytes`-sized device buffer streamed `iters` times with
 * the traversal kernel's 256-bit read-only loads: L2 for buffers that fit it, HBM beyond. */
int nnbvh_read_bandwidth_probe(nnbvh_engine *e, uint64_t bytes, uint32_t iters, float *gbs);

/* Diagnostic: lane-occupancy counters of the traversal kernel, accumulated since the
 * last call, when the engine was created with NNBVH_STEP_STATS=1 in the environment
 * (a counting instantiation of the kernel runs then).  out8: phase-A iterations, lanes
 * at a node, lanes parked at a leaf, lanes without a ray, phase-B passes, lanes parked
 * there, lanes that ran a triangle test, refills (all summed over warps). */
int nnbvh_debug_step_stats(nnbvh_engine *e, uint64_t *out8);

/* Replaces the reference's pointer graph (Primitive -> SimplePrimitive ->
 * Triangle -> TriangleMesh, cpu/primitive.cpp:96-109, shapes.cpp:320-327) by
 * device arrays.  nodes: the flattened tree; orderedTriVerts9: nPrims*9 floats in
 * ORDERED (leaf) order.  Host or device pointers.  nNodes == 0 uploads the empty
 * scene (`nodes == nullptr` in the reference: every ray misses). */
int nnbvh_upload_scene(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                       const float *orderedTriVerts9, uint32_t nPrims);

/* nnbvh_upload_scene for a primitive list that mixes triangles and spheres (the killeroos scene's
 * spherical light, crown's environment shell).  orderedSphereIndex[i] = -1 when ordered primitive
 * i is a triangle (its 9 floats are read from orderedTriVerts9 as usual), else the index of its
 * sphere in `spheres` (its 9 floats are ignored).  A hit on a sphere is reported as
 * {prim, t = tHit, b0,b1,b2 = pObj.xyz}: QuadricIntersection (shapes.h:100-104) without phi,
 * which is atan2(pObj.y, pObj.x) (+2*Pi when negative) on the host;
 * Sphere::InteractionFromIntersection (shapes.h:239-300) takes it from there.  The per-ray
 * counter triTests counts triangle tests only, as the reference's nTriTests does. */
int nnbvh_upload_scene_mixed(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                             const float *orderedTriVerts9, const int32_t *orderedSphereIndex,
                             uint32_t nPrims, const nnbvh_sphere *spheres, uint32_t nSpheres);

/* ---- alpha-tested triangles: GeometricPrimitive with an alpha texture (cpu/primitive.cpp:50-84) ------
 * orderedAlpha[i]: the CONSTANT alpha of ordered primitive i (FloatConstantTexture; >= 1 = no texture).
 * A hit is ignored when alpha <= 0 or HashFloat(ray.o, ray.d) > alpha (util/hash.h:98-112), and the same
 * primitive is asked again with si->intr.SpawnRay(ray.d) and tMax - tHit, as the reference does;
 * nnbvh_intersect_any on such a scene is Intersect(...).has_value().  Image / procedural alpha textures
 * are outside this path.  Such scenes run the product kernel with the alpha test in its leaf phase
 * (NNBVH_INST_STEP=0: a one-thread-per-ray kernel); per-ray counters are unavailable. */
int nnbvh_upload_scene_alpha(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                             const float *orderedTriVerts9, const float *orderedAlpha, uint32_t nPrims);

/* ---- instances: TransformedPrimitive over per-object BVHAggregates ---------------------------
 * (cpu/primitive.cpp:112-130; pbrt creates them for ObjectInstance, scene.cpp:1541-1571).  An object
 * is a BVHAggregate over triangles in ITS OWN space (tree + ordered triangles, as nnbvh_upload_scene
 * takes them); an instance places one object with renderFromPrimitive.  The top-level primitive list
 * mixes triangles and instances: orderedInstanceIndex[i] = -1 for a triangle, else the instance.
 * Each ray is moved into the instance's space exactly as Transform::ApplyInverse(Ray, &tMax) does
 * (interval error bound of the origin, origin pushed along d, tMax shortened) and the object's tree is
 * traversed there.  A hit on an instance is {prim = the instance's top-level slot, t and b0,b1,b2 of
 * the inner triangle in the instance ray's parametrisation} -- what TransformedPrimitive::Intersect
 * returns -- plus, from nnbvh_intersect_closest_instanced, the ordered primitive inside the object.
 * Such scenes run the product kernel (StepKernel with instance slots: the objects' nodes and triangles are
 * appended to the top level's device arrays and the second level is entered as a pushed sub-state);
 * per-ray counters, NNBVH_INST_STEP=0 and scenes whose combined stack depth exceeds 64 entries run the
 * one-thread-per-ray kernel.  Spheres and alpha-tested triangles beside instances:
 * nnbvh_upload_scene_general. */
typedef struct nnbvh_object {
    const nnbvh_node *nodes;
    uint32_t nNodes;
    const float *orderedTriVerts9;
    uint32_t nPrims;
} nnbvh_object;
typedef struct nnbvh_instance {
    float renderFromPrimitive[12]; /* rows 0..2 of the 4x4, row 3 = 0 0 0 1 */
    float primitiveFromRender[12]; /* its inverse, as Transform caches it (mInv) */
    uint32_t object;
    uint32_t reserved[3];
} nnbvh_instance;
/* TransformedPrimitive::Bounds() given the object's root bounds (its nodes[0].bounds). */
int nnbvh_instance_bounds(const nnbvh_instance *in, const float *objectBounds6, float *bounds6);
int nnbvh_upload_scene_instanced(nnbvh_engine *e, const nnbvh_node *nodes, uint32_t nNodes,
                                 const float *orderedTriVerts9, const int32_t *orderedInstanceIndex,
                                 uint32_t nPrims, const nnbvh_instance *instances, uint32_t nInstances,
                                 const nnbvh_object *objects, uint32_t nObjects);
/* nnbvh_intersect_closest on an instanced scene, also returning innerPrim[i] = ordered primitive
 * inside the hit object (NNBVH_MISS for a top-level triangle or a miss).  innerPrim may be NULL. */
int nnbvh_intersect_closest_instanced(nnbvh_engine *e, const nnbvh_ray *rays, uint64_t n,
                                      nnbvh_hit *hits, uint32_t *innerPrim);
/* The same over the wavefront queues' seven planes (wavefront/workitems.soa; tMax NULL = Infinity). */
int nnbvh_intersect_closest_instanced_soa(nnbvh_engine *e, const float *ox, const float *oy,
                                          const float *oz, const float *dx, const float *dy,
                                          const float *dz, const float *tMax, uint64_t n,
                                          nnbvh_hit *hits, uint32_t *innerPrim);

/* ---- general scenes: any mix of primitive kinds in ONE list -----------------------------------
 * What pbrt hands to one BVHAggregate (scene.cpp:1352-1591): SimplePrimitive / GeometricPrimitive over
 * Triangle or Sphere, GeometricPrimitive with a constant alpha texture over Triangle, TransformedPrimitive
 * over per-object aggregates -- all in the same std::vector<Primitive> (cpu/primitive.cpp:50-130).
 * Per ordered primitive i: orderedSphereIndex[i] >= 0 -> sphere, orderedInstanceIndex[i] >= 0 -> instance
 * (never both), else the triangle orderedTriVerts9[9 i ..] with orderedAlpha[i] (>= 1 or a NULL array: no
 * alpha test).  Index arrays of kinds that do not occur may be NULL.  Hit records are those of the
 * specialised uploads: triangle {prim, t, b}, sphere {prim, tHit, pObj}, instance {prim = top-level slot,
 * inner triangle's t, b} (+ innerPrim from nnbvh_intersect_closest_instanced).  A description with at most
 * one extra kind is routed to that kind's own upload and kernel; a real mix runs the general
 * instantiation of the product kernel (sphere leaves + instance slots + alpha test), and per-ray
 * counters are unavailable for it. */
typedef struct nnbvh_scene_desc {
    const nnbvh_node *nodes;
    uint32_t nNodes;
    const float *orderedTriVerts9;
    uint32_t nPrims;
    const int32_t *orderedSphereIndex;
    const nnbvh_sphere *spheres;
    uint32_t nSpheres;
    const int32_t *orderedInstanceIndex;
    const nnbvh_instance *instances;
    uint32_t nInstances;
    const nnbvh_object *objects;
    uint32_t nObjects;
    const float *orderedAlpha;
} nnbvh_scene_desc;
int nnbvh_upload_scene_general(nnbvh_engine *e, const nnbvh_scene_desc *d);

/* BVHAggregate::Intersect (cpu/aggregates.cpp:529-579) for n rays.  hits[i] is final when the call's work on
 * the stream has completed; while it runs, a ray's slot may hold the record of a hit that a closer one later
 * replaces (the kernel writes a record when a hit is accepted). */
int nnbvh_intersect_closest(nnbvh_engine *e, const nnbvh_ray *rays, uint64_t n,
                            nnbvh_hit *hits);

/* BVHAggregate::IntersectP (cpu/aggregates.cpp:581-624) for n rays: occluded[i] = 0/1 (all n bytes are written:
 * zeroed first, then the occluded rays' 1s). */
int nnbvh_intersect_any(nnbvh_engine *e, const nnbvh_ray *rays, uint64_t n,
                        uint8_t *occluded);

/* nnbvh_intersect_closest for callers that cannot allocate: code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * Same kernels and bit-identical results as nnbvh_intersect_closest
 * with device pointers, but the per-launch scratch (ray cursors, the list of rays deferred to the
 * exact slab test) lives in a file-scope `static __device__` array inside the library -- 256 B +
 * 4 B x NNBVH_NOALLOC_MAX_RAYS per GPU that loads the library, held until the process exits -- so
 * the call makes no cudaMalloc / cudaFree or any allocating library call: one cudaMemsetAsync and
 * the two traversal launches on the engine's stream, asynchronously.
 *   - raysDev / hitsDev must be DEVICE pointers (host buffers would need staging memory);
 *   - n > NNBVH_NOALLOC_MAX_RAYS fails with NNBVH_ERR_INVALID (split the batch, or use
 *     nnbvh_intersect_closest); no ray-ordering pre-pass runs (it needs O(n) buffers);
 *   - triangle and sphere scenes only (alpha / instanced scenes: NNBVH_ERR_INVALID);
 *   - every call on a device shares the one array: calls that OVERLAP on the GPU (different
 *     streams, engines or threads without ordering between them) race; calls ordered on one
 *     stream are safe, the library clears what must start at zero at the start of each call;
 *   - the array's address is resolved and the kernels are loaded by nnbvh_create. */
#define NNBVH_NOALLOC_MAX_RAYS (1u << 20)
int nnbvh_intersect_closest_noalloc(nnbvh_engine *e, const nnbvh_ray *raysDev, uint64_t n,
                                    nnbvh_hit *hitsDev);

/* SoA forms for the wavefront queues (SOA<RayWorkItem>, wavefront/workitems.soa;
 * wavefront/aggregate.cpp:34-68).  tMax may be NULL (= Infinity). */
int nnbvh_intersect_closest_soa(nnbvh_engine *e, const float *ox, const float *oy,
                                const float *oz, const float *dx, const float *dy,
                                const float *dz, const float *tMax, uint64_t n,
                                nnbvh_hit *hits);
int nnbvh_intersect_any_soa(nnbvh_engine *e, const float *ox, const float *oy,
                            const float *oz, const float *dx, const float *dy,
                            const float *dz, const float *tMax, uint64_t n,
                            uint8_t *occluded);

/* Same results plus the reference's per-ray statistics: nodesVisited counts as
 * `++nodesVisited` does (cpu/aggregates.cpp:541,592), triTests as `++nTriTests`
 * (shapes.cpp:322,343).  Runs the BASIC kernel.  anyHit != 0 -> IntersectP
 * (hits may be NULL, occluded filled); else Intersect (occluded may be NULL). */
int nnbvh_intersect_counters(nnbvh_engine *e, const nnbvh_ray *rays, uint64_t n,
                             int32_t anyHit, nnbvh_hit *hits, uint8_t *occluded,
                             uint32_t *nodesVisited, uint32_t *triTests);

/* A second engine on the same device that SHARES src's uploaded scene (no copy; the device arrays live
 * until the last engine holding them is destroyed) and has its own stream, scratch and host-pointer
 * pipeline.  One engine serialises its batches; an integrator that has independent batches in flight --
 * the shadow rays of bounce k and the closest-hit rays of bounce k+1 -- gives each its own clone and
 * stream: the second batch's ray-ordering pre-pass and first CTAs fill the SMs the first batch's
 * persistent CTAs leave as it drains (+5 % on the bench workload, tools/overlap_probe.py).  A later upload
 * into either engine leaves the other with the old scene. */
int nnbvh_clone(nnbvh_engine *src, nnbvh_engine **out);

/* Device time (CUDA events on the engine's stream) of the traversal kernel(s)
 * of the most recent intersect call, and kernels launched so far. */
int nnbvh_last_kernel_ms(nnbvh_engine *e, float *ms);
int nnbvh_launch_count(nnbvh_engine *e, uint64_t *launches);

/* IntersectTriangle (shapes.cpp:172-273) and Bounds3f::IntersectP
 * (util/vecmath.h:1573-1608) evaluated ON THE DEVICE for n independent
 * (ray, triangle) / (ray, box) pairs -- known-answer hooks for the parity
 * tests.  out4 = {t,b0,b1,b2}; hit[i] = 0/1.  Host pointers. */
int nnbvh_test_triangles(nnbvh_engine *e, const nnbvh_ray *rays, const float *tris9,
                         uint64_t n, uint8_t *hit, float *out4);
int nnbvh_test_boxes(nnbvh_engine *e, const nnbvh_ray *rays, const float *boxes6,
                     uint64_t n, uint8_t *hit);
/* Sphere::BasicIntersect (shapes.h:147-232) on n independent (ray, sphere) pairs.
 * out4[i] = {tHit, pObj.x, pObj.y, pObj.z} where hit[i] != 0. */
int nnbvh_test_spheres(nnbvh_engine *e, const nnbvh_ray *rays, const nnbvh_sphere *spheres,
                       uint64_t n, uint8_t *hit, float *out4);
/* The device's std::atan2(float, float) -- the host libm's algorithm restated, which the phimax
 * clipping test of a partial sphere compares (shapes.h:196-203) -- on n (y, x) pairs.  Host pointers. */
int nnbvh_test_atan2(nnbvh_engine *e, const float *y, const float *x, uint64_t n, float *out);

const char *nnbvh_last_error(void);
const char *nnbvh_version(void);

#ifdef __cplusplus
}
#endif
#endif /* NNBVH_H */
