// rt_kernels.cuh -- kernel parameter blocks and host-side launchers (defined in the .cu files).
#pragma once
#include "rt_scene.cuh"

namespace rtb {

#define RT_STAT_SLOTS 64           // counters are spread over 64 slots to keep atomics apart
enum {                             // per-slot counter index
    ST_SHADOW = 0, ST_REFLECT, ST_REFRACT, ST_NAN, ST_TEST_TRI, ST_TEST_QUAD, ST_TEST_PLANE, ST_BVH_NODES,
    ST_BVH_OVERFLOW, ST_TEST_TRI_SHADOW, ST_BVH_NODES_SHADOW, ST_TEST_TRI_EXEC,
    ST_PEER_TIMEOUT,               // CTAs that gave up waiting for rank 0's acknowledgement (peer frame): the frame may be torn
    ST_SHADOW_FIXUP,               // shadow queries whose filter product was redone in insertion order (rt_cullscan.cu)
    ST_BVH_OUT_OF_RADIUS,          // LBVH rays that started beyond the radius the boxes are padded for (took the widened walk)
    ST_COUNT
};

// One wavefront level of the scene-resident fused path (trace + shade + shadow + spawn).
struct TailLevel {                 // one deep level of a tail-mode launch (see LevelParams::tail)
    const float4* in_od0;
    const float4* in_od1;
    float4* in_colour;
    float4* out_od0;
    float4* out_od1;
    int4*   links;
    int     level, in_cap, spawn;   // spawn: 0 = the deepest allowed level
    unsigned int out_cap;           // 0 = a probe: may only report that it wanted children
};

struct LevelParams {
    FrameParams f;
    const unsigned char* blob;     // resident blob (global)
    ResidentLayout rl;
    int   level;                   // 0 = primary rays generated from the lattice
    // level 0: chunk of the local sample lattice
    int   chunk_ly0, chunk_rows, local_LH;
    float* c0;                     // [chunk_rows][LW][3] sample colours
    // level >= 1: input queue
    const float4* in_od0;
    const float4* in_od1;
    float4* in_colour;
    int   in_count;
    // children
    float4* out_od0;
    float4* out_od1;
    unsigned int* out_count;
    int4* links;
    unsigned int* link_count;
    unsigned long long* stats;     // [RT_STAT_SLOTS][ST_COUNT]
    // direct resolve (level 0, spp 1, no secondary rays): the level kernel writes the frame itself
    int     cull;                  // warp-vote culling of bounded quadrics / triangles (0 with RTB200_F_EXACT_FP)
    int     pre_table;             // set by the launcher: shared memory holds room for the primary-ray culling table
    int     direct;
    float*  fb_f32;                // [local rows][W][3]
    uchar4* fb_rgba8;              // [local rows][W]
    // level loop without host round trips (rt_api.cu, render_resident_async): the number of input rays is
    // read from the counter the previous level wrote, and appends are bounded by the capacity the host
    // guessed -- a kernel that runs out of room raises *overflow and the host redoes the frame the slow way
    const unsigned int* in_count_dev;   // null: in_count is exact
    int   in_cap;                       // entries the input queue can hold
    int   out_guarded;                  // 1: appends are checked against out_cap
    unsigned int out_cap;
    unsigned int* overflow;
    // tail mode (level >= 1 kernels, ONE CTA): the launch walks tail_count consecutive deep levels itself, a block
    // barrier between them, taking each level's queues from this table in global memory -- a few hundred rays per
    // level are not worth a launch each (render_resident_async)
    const struct TailLevel* tail;
    int   tail_count;
    // programmatic dependent launch: 1 = this kernel is part of a chain enqueued in one go; it lets its
    // successor start early (the successor's prologue overlaps this kernel's tail) and waits for its
    // predecessor before touching anything the predecessor wrote.  2 = same, and the launch itself carries
    // the attribute (every kernel of the chain but the first).
    int   pdl;
    // several GPUs, peer frame (rt_comm.cu): the kernel that stored this rank's rows into rank 0's
    // memory also tells rank 0 so -- the last CTA to finish writes peer_seq to peer_flag over NVLink
    unsigned int* peer_flag;       // null = no signal from this launch
    unsigned int* peer_done;       // local CTA counter (zero between launches)
    unsigned int  peer_seq;
    // ... and may first have to wait until rank 0 has let go of the frame half it is about to overwrite
    const unsigned int* peer_ack;  // null = nothing to wait for (or the stream already waited)
    unsigned int  peer_ack_min;
};

struct CombineParams {
    const int4* links;
    int   n_links;
    float*  parent_c0;             // level-0 parents: 3 floats per sample
    float4* parent_colour;         // level >= 1 parents
    const float4* child_colour;
    const MatRec* mat;
    int   parent_is_level0;
    const unsigned int* n_links_dev;    // non-null: the link count lives on the device (n_links = capacity)
    int   pdl;                          // see LevelParams::pdl
};

// fold of SEVERAL deep levels by one CTA (render_resident_async): levels last, last - 1, ..., first (first >= 1),
// a block barrier between them.  Deep levels hold a few hundred links; one launch instead of one per level.
#define RT_TAIL_FOLD_MAX 32
struct TailFoldParams {
    int first, last;
    const int4*  links[RT_TAIL_FOLD_MAX];      // [l - first]
    float4*      colour[RT_TAIL_FOLD_MAX + 1]; // [l - first] parents of level l, [l - first + 1] their children
    int          cap[RT_TAIL_FOLD_MAX];
    const unsigned int* counters;              // [2 * l + 1] = links at level l
    const MatRec* mat;
    int pdl;
};

struct ResolveParams {
    FrameParams f;
    const float* c0;               // [chunk_rows][LW][3]
    int   chunk_py0, chunk_prows;  // local pixel rows covered by the chunk
    float*  fb_f32;                // [local rows][W][3] or null
    uchar4* fb_rgba8;              // [local rows][W]    or null
    unsigned long long* stats;
    int   pdl;                     // see LevelParams::pdl
};

// scene packing on the device (rt_pack.cu): raw caller arrays -> unified list + tile arrays
struct PackParams {
    int T, Tpad, Q, P;
    const float* tri_points;       // [T][9]   raw, as the caller passed them
    const float* tri_normal;       // [T][3]
    const int*   tri_material;     // [T]
    const int*   tri_order;        // [T]
    const float4* quad_r3;         // [Q] centred form for the culling votes (computed on the host in double)
    SurfRec* surf;                 // out: unified list (every slot is written by exactly one thread)
    TriRec*  tri;                  // out: [Tpad]
    float4*  trn;                  // out: [Tpad]
    const QuadRec*  quad;          // [Q] already in place (uploaded packed)
    const PlaneRec* plane;         // [P] already in place
};
cudaError_t launch_pack_scene(const PackParams& p, cudaStream_t st);

// launchers (rt_resident.cu)
size_t resident_smem_bytes(const ResidentLayout& rl);
cudaError_t resident_prepare(size_t max_smem_bytes);
cudaError_t launch_resident_level(const LevelParams& p, bool exact, int sm_count, cudaStream_t st);
cudaError_t launch_combine(const CombineParams& p, bool exact, cudaStream_t st);
cudaError_t launch_tail_fold(const TailFoldParams& p, bool exact, cudaStream_t st);
cudaError_t launch_resolve(const ResolveParams& p, cudaStream_t st);

// ray-level queries over the unified list in global memory (rt_resident.cu)
cudaError_t launch_query_nearest(const DeviceScene& s, int n, const float* o3, const float* d3,
                                 int* out_order, float* out_t, float* out_n3, bool exact, bool cull, cudaStream_t st);
cudaError_t launch_query_shadow(const DeviceScene& s, int n, const float* p3, const float* l3,
                                float* out3, bool exact, bool cull, cudaStream_t st);

// FFMA peak micro-benchmark (rt_peak.cu): returns FFMA-chain flop count and elapsed ms
cudaError_t run_fp32_peak(int sm_count, double* tflops, cudaStream_t st, double* detail = nullptr);

} // namespace rtb
