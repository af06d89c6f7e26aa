/* rtb200.h -- C ABI of librtb200.so, the B200-native (sm_100a CUDA) render path that
 * replaces the reference's GLSL compute-shader dispatch.
 *
 * What it replaces (file:line relative to Rodyll/Ray-Tracing-GPU, Projet-INF8702/):
 *   - CNuanceurCalculProg (NuanceurCalculProg.h:73-132): create / compile / activate one
 *     compute program                                   -> rtb200_create / rtb200_destroy
 *   - the two std140 UBO uploads of CScene::LancerRayons (Scene.cpp:704-783 "General",
 *     Scene.cpp:799-1269 "SceneData": glBufferData + glBindBufferBase)
 *                                                        -> rtb200_set_scene (+ rtb200_frame)
 *   - glBindImageTexture + glDispatchCompute(W/16,H/16,1) + glMemoryBarrier
 *     (Scene.cpp:697-698, 1307-1310) and the kernel rayTracing.glsl:128-509
 *                                                        -> rtb200_render
 * The plug-in point stays where the reference has it: the `if (CVar::g_ComputerShadersON)`
 * branch of CScene::LancerRayons() (Scene.cpp:692).  The results are those of the
 * reference's C++ CPU renderer (Scene.cpp:1535-1563, 1705-1861), with the reflection /
 * refraction block of Scene.cpp:1779-1823 enabled and bounded by max_bounces.
 *
 * Conventions
 *   - plain C, plain pointers and sizes; every entry point returns an int status
 *     (0 = RTB200_OK, negative = RTB200_E_*); nothing exits, nothing throws.
 *   - the caller owns every host array it passes; the library copies during the call and
 *     keeps no host pointer.  The context owns all device memory, streams, events and the
 *     NCCL communicator.
 *   - a context is bound to ONE GPU and is not thread-safe (one caller thread), like the
 *     GL context it replaces.  Multi-GPU = one context per GPU (one process or thread per
 *     GPU) joined with rtb200_comm_init; image row-bands are sharded across the ranks and
 *     the framebuffer is gathered to rank 0 with NCCL inside rtb200_render.
 *   - all geometry is FINAL: ISurface::Pretraitement (Triangle.cpp:108, Plan.cpp:101,
 *     Quadrique.cpp:110) has already been applied by the host scene classes.
 *   - images: row 0 is the BOTTOM scanline (Scene.cpp:1556: index (x + y*W)), float RGB is
 *     unclamped, RGBA8 is floor(clamp(c,0,1)*255+0.5) with A = 255 and NaN -> 0.
 *   - arithmetic: by default every value that is kept (hit distance, normal, colour) is
 *     produced by the reference's own binary32 rounding sequence (one rounding per operation,
 *     no contraction), so frames are bit-identical to the CPU renderer except where powf
 *     enters (<= 4 ulp of the Phong term); fused multiply-adds are used only inside
 *     conservative culling tests whose survivors are re-evaluated in the reference sequence.
 *   - there is NO CPU fallback: without a usable CUDA device every compute entry point
 *     fails with RTB200_E_CUDA.
 */
#ifndef RTB200_H
#define RTB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB200_ABI_VERSION 1

enum {
    RTB200_OK            =  0,
    RTB200_E_ARG         = -1,   /* bad argument / inconsistent sizes                        */
    RTB200_E_CUDA        = -2,   /* CUDA runtime error (text: rtb200_last_error)             */
    RTB200_E_NCCL        = -3,   /* NCCL error or NCCL library not loadable                   */
    RTB200_E_OOM         = -4,   /* device or pinned-host allocation failed                   */
    RTB200_E_CAPACITY    = -5,   /* secondary-ray queues overflowed even at the smallest chunk */
    RTB200_E_STATE       = -6    /* call order (e.g. render before set_scene, bvh not built)  */
};

/* rtb200_frame.flags */
enum {
    RTB200_F_USE_BVH      = 1u << 0,  /* triangles through the GPU-built LBVH (rtb200_build_bvh)        */
    RTB200_F_NO_SECONDARY = 1u << 1,  /* as-shipped reference: no reflection / refraction rays           */
    RTB200_F_EXACT_FP     = 1u << 2,  /* reference rounding sequence everywhere AND no conservative culling */
                                      /* filters: every test is the reference's own arithmetic            */
    RTB200_F_FORCE_STREAM = 1u << 3,  /* always use the tiled (streamed) brute-force kernels              */
    RTB200_F_FORCE_RESIDENT = 1u << 4,/* always use the scene-resident fused kernel (fails if too large)  */
    RTB200_F_FUSED_FP     = 1u << 5,  /* contract every a*b+c into one FMA (fastest; NOT the reference's   */
                                      /* rounding: decisions on acne / edge pixels may flip) */
    RTB200_F_GATHER_F32   = 1u << 6,  /* multi-GPU: also gather the float RGB frame (default: RGBA8 only)   */
    RTB200_F_HOST_BANDS   = 1u << 7   /* several ranks: the host pointers of rtb200_render are the FULL frame on */
                                      /* EVERY rank (one buffer shared by the ranks: threads of a process, or a  */
                                      /* shared-memory segment between processes) and each rank copies its own    */
                                      /* bands into it over its own PCIe link; nothing is gathered on the device  */
};
#define RTB200_MAX_BOUNCES 4096       /* rtb200_frame.max_bounces above this is refused (reference default: 20)  */

/* Material block of ISurface (ISurface.h:24-41; defaults ISurface.cpp:15-25). 10 floats. */
typedef struct rtb200_material {
    float color[3];      /* m_Couleur (already /255 when it came from a scene file)          */
    float ambient;       /* m_CoeffAmbient   (0.2)                                           */
    float diffuse;       /* m_CoeffDiffus    (0.8)                                           */
    float specular;      /* m_CoeffSpeculaire                                                */
    float shininess;     /* m_CoeffShininess                                                 */
    float reflect;       /* m_CoeffReflexion                                                 */
    float refract;       /* m_CoeffRefraction                                                */
    float ior;           /* m_IndiceDeRefraction                                             */
} rtb200_material;

/* CLumiere (Lumiere.h:19-48). 7 floats. */
typedef struct rtb200_light {
    float position[3];
    float color[3];
    float intensity;
} rtb200_light;

/* Scene = what the "SceneData" UBO carried (rayTracing.glsl:97-107), without its caps of
 * 10 per type, as structure-of-arrays.  *_order[i] is the insertion index of the primitive
 * in CScene::m_Surfaces (Scene.cpp:107): ties in distance go to the lowest order, exactly as
 * the linear scan of Scene.cpp:1710-1715 resolves them.  Orders must be a permutation of
 * 0..T+Q+P-1 and increasing within each array. */
typedef struct rtb200_scene {
    int32_t n_triangles, n_quadrics, n_planes, n_lights, n_materials;
    const float*   tri_points;     /* [T][3][3]  m_Pts after Pretraitement (Triangle.h:20)           */
    const float*   tri_normal;     /* [T][3]     m_Normale (Triangle.cpp:199-204)                    */
    const int32_t* tri_material;   /* [T]                                                             */
    const int32_t* tri_order;      /* [T]                                                             */
    const float*   quad_coef;      /* [Q][10]    A B C | mixed yz xz xy | linear x y z | K  (Quadrique.h:20-29) */
    const int32_t* quad_material;  /* [Q] */
    const int32_t* quad_order;     /* [Q] */
    const float*   plane_coef;     /* [P][4]     normal xyz, constant (Plan.h:11-12)                 */
    const int32_t* plane_material; /* [P] */
    const int32_t* plane_order;    /* [P] */
    const rtb200_material* materials; /* [M] */
    const rtb200_light*    lights;    /* [L] */
} rtb200_scene;

/* Frame = what the "General" UBO carried (rayTracing.glsl:79-91) plus the CScene knobs the
 * CPU renderer reads (Scene.cpp:68-70, 162, 180, 197, 214). */
typedef struct rtb200_frame {
    int32_t  width, height;        /* pixels; no multiple-of-16 / power-of-two restriction           */
    int32_t  spp_grid;             /* k: k*k regular sub-pixel grid, sample (i,j) of pixel (X,Y) is   */
                                   /* the reference's primary ray at (k*X+i, k*Y+j) of a kW x kH      */
                                   /* image; pixel = mean of the k*k float RGB results.  1 = reference */
    int32_t  max_bounces;          /* m_NbRebondsMax (reference default 20)                           */
    float    min_energy;           /* m_EnergieMinRayon (0.01)                                        */
    float    scene_ior;            /* m_IndiceRefractionScene (1.0)                                   */
    float    background[3];        /* m_CouleurArrierePlan                                            */
    float    cam_pos[3];           /* m_Camera.Position                                               */
    float    cam_u[3], cam_v[3], cam_n[3]; /* rows 0,1,2 of m_Camera.Orientation (Scene.cpp:656-659)  */
    float    half_w, half_h;       /* Scene.cpp:676-677                                               */
    uint32_t flags;                /* RTB200_F_*                                                      */
    int32_t  band_rows;            /* multi-GPU: height of the interleaved row bands (0 = default 16) */
} rtb200_frame;

/* Device-side measurements of the last rtb200_render on THIS rank. */
typedef struct rtb200_stats {
    double   ms_render;            /* CUDA-event time, first kernel .. last kernel of the frame       */
    double   ms_gather;            /* NCCL gather + de-interleave (0 without a communicator)          */
    double   ms_h2d, ms_d2h;       /* copies done inside the call (0 when none were requested)        */
    uint64_t rays_primary, rays_shadow, rays_reflect, rays_refract;
    uint64_t tests_triangle, tests_quadric, tests_plane;   /* ray-primitive tests entered             */
    uint64_t bvh_nodes_visited;
    uint64_t nan_pixels;           /* NaN channels quantised to 0                                     */
    uint32_t kernel_launches;      /* kernels launched by the call                                    */
    uint32_t depth_levels;         /* wavefront levels that held at least one ray                     */
    uint32_t path;                 /* 0 = scene-resident fused kernels, 1 = streamed tiled kernels, 2 = LBVH */
    uint32_t chunks;               /* image chunks the frame was split into                           */
    double   ms_trace_nearest, ms_trace_shadow, ms_shade;  /* streamed / LBVH path: per-stage device  */
                                   /* totals over all levels and chunks (resident path: the fused kernels) */
    double   ms_primary_kernel;    /* the level-0 nearest-hit (or fused resident) kernel of the first chunk */
    uint64_t shadow_tests_triangle;/* part of tests_triangle entered by shadow-filter queries              */
    uint64_t shadow_bvh_nodes;     /* part of bvh_nodes_visited entered by shadow-filter queries           */
    uint64_t level_loop_redone;    /* 1: the frame's levels were first enqueued without host round trips,   */
                                   /* a guessed queue capacity proved too small and the frame was redone    */
                                   /* with exact capacities (same pixels, slower)                           */
    uint64_t tests_triangle_executed; /* streamed path: (ray, triangle) tests actually entered after the warp-level */
                                   /* bundle culling (tests_triangle counts what the reference's scan performs)   */
    uint64_t shadow_fixups;        /* streamed path: shadow queries whose filter product (>= 3 transparent occluders */
                                   /* met out of insertion order) was redone in the order of Scene.cpp:1856          */
    uint64_t bvh_out_of_radius;    /* LBVH: rays that started farther from the mesh than the radius its boxes are    */
                                   /* padded for and were walked with a pad widened for their own origin             */
} rtb200_stats;

typedef struct rtb200_ctx rtb200_ctx;

/* --- lifetime --------------------------------------------------------------------------- */
int  rtb200_abi_version(void);
/* Creates a context on CUDA device `device_id` (the replacement for constructing
 * CNuanceurCalculProg + compilerEtLier, NuanceurCalculProg.cpp:57,280). */
int  rtb200_create(int device_id, rtb200_ctx** out);
void rtb200_destroy(rtb200_ctx* ctx);
const char* rtb200_error_string(int status);
/* Text of the last CUDA/NCCL failure seen by this context ("" if none). */
const char* rtb200_last_error(const rtb200_ctx* ctx);

/* --- scene -------------------------------------------------------------------------------- */
/* Validates (see rtb200_validate_scene), packs to the device layout and uploads: small scenes are packed by the
 * host into one pinned staging buffer (one async copy), large meshes are uploaded raw and packed by a kernel. */
int  rtb200_set_scene(rtb200_ctx* ctx, const rtb200_scene* scene);
/* The checks rtb200_set_scene applies, without a context or a GPU: counts, null arrays, material indices,
 * and the three order arrays (each strictly increasing, together a permutation of 0..N-1 = the insertion
 * indices of CScene::m_Surfaces, Scene.cpp:107).  *why (may be NULL) receives a static string. */
int  rtb200_validate_scene(const rtb200_scene* scene, const char** why);
/* Optional: GPU-built LBVH over the triangles of the current scene (Morton codes, radix sort,
 * Karras hierarchy, bottom-up refit).  Planes and quadrics stay in the brute-force side list. */
int  rtb200_build_bvh(rtb200_ctx* ctx);
/* Facts about the LBVH of the current scene: internal nodes, triangles, device build time (ms),
 * the origin radius its boxes are padded for and bytes per node. */
int  rtb200_bvh_info(rtb200_ctx* ctx, int* n_nodes, int* n_triangles, double* build_ms,
                     double* pad_radius, double* node_bytes);

/* --- render ------------------------------------------------------------------------------- */
/* Renders one frame.  out_rgb_f32 (W*H*3 floats) and out_rgba8 (W*H*4 bytes) are HOST
 * pointers, either may be NULL (nothing is copied back for a NULL pointer; the frame stays
 * on the device and can be fetched later with rtb200_read_frame).  With a communicator the
 * frame is sharded over the ranks in interleaved row bands and gathered on rank 0; the
 * output pointers are ignored on the other ranks.  Synchronous: returns when the frame is
 * complete (and copied, if requested). */
int  rtb200_render(rtb200_ctx* ctx, const rtb200_frame* frame,
                   float* out_rgb_f32, uint8_t* out_rgba8, rtb200_stats* stats);
int  rtb200_read_frame(rtb200_ctx* ctx, float* out_rgb_f32, uint8_t* out_rgba8);
/* A camera path as ONE pipelined submission (the reference renders a single frame per process,
 * Main.cpp:181; repeated frames with the scene resident are SURVEY.md 8f rank 2).  All frames share
 * width, height, spp_grid, flags and band_rows.  out_rgba8 (HOST, n_frames * W*H*4 bytes, frame i at
 * offset i * W*H*4) and stats ([n_frames]) may be NULL.  Frames whose level-0 kernel is the whole
 * frame (scene-resident path, no secondary rays, spp_grid 1) are enqueued back to back with no host
 * round trip between them, read back on a second stream while the next frame renders, and -- with a
 * communicator and peer access -- signalled with one asynchronous all-reduce per frame; any other
 * sequence is rendered frame by frame through rtb200_render (identical results).  With several
 * ranks out_rgba8 must be NULL for the pipelined form; the last frame stays on rank 0
 * (rtb200_read_frame).  Per-frame ms_render of a pipelined sequence = sequence time / n_frames. */
int  rtb200_render_sequence(rtb200_ctx* ctx, const rtb200_frame* frames, int n_frames,
                            uint8_t* out_rgba8, rtb200_stats* stats);

/* Ray-level entry points used by the parity tests (the per-ray oracle of SURVEY.md 8c has
 * the same shape): n rays in, nearest hit (insertion order index or -1, distance, normal)
 * or shadow-filter RGB out.  Host pointers.  `flags` accepts RTB200_F_USE_BVH / EXACT_FP /
 * FORCE_*. */
int  rtb200_trace_nearest(rtb200_ctx* ctx, int n, const float* origin3, const float* dir3,
                          uint32_t flags, int32_t* out_order, float* out_t, float* out_normal3);
int  rtb200_trace_shadow(rtb200_ctx* ctx, int n, const float* point3, const float* light3,
                         uint32_t flags, float* out_filter3);

/* rtb200_trace_nearest WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same results as rtb200_trace_nearest, but the device-side staging of the rays and
 * of the answers lives in a static __device__ array inside the library (RTB200_STATIC_QUERY_MAX rays x 44 bytes =
 * 2.75 MB, held by every GPU that loads the library until the process exits); the call makes no cudaMalloc / cudaFree
 * and calls no allocating library routine.  Limits: n <= RTB200_STATIC_QUERY_MAX (more: RTB200_E_ARG, split the batch);
 * only scenes that take the scene-resident query path (packed scene <= 64 KB, or RTB200_F_FORCE_RESIDENT; not
 * RTB200_F_FORCE_STREAM / RTB200_F_USE_BVH, whose kernels keep scratch structures of their own: RTB200_E_ARG).
 * All calls on one GPU share the array (it keeps the last call's contents; nothing in it needs clearing): calls that
 * overlap on the same GPU -- two contexts on one device driven from two threads -- race and must be serialised. */
#define RTB200_STATIC_QUERY_MAX 65536
int  rtb200_trace_nearest_static(rtb200_ctx* ctx, int n, const float* origin3, const float* dir3,
                                 uint32_t flags, int32_t* out_order, float* out_t, float* out_normal3);

/* --- multi-GPU ---------------------------------------------------------------------------- */
#define RTB200_UNIQUE_ID_BYTES 128
/* Rank 0 obtains an id, distributes the bytes by any means (torch.distributed, a file, MPI),
 * then every rank calls rtb200_comm_init with the same id. */
int  rtb200_comm_unique_id(uint8_t id[RTB200_UNIQUE_ID_BYTES]);
int  rtb200_comm_init(rtb200_ctx* ctx, int rank, int n_ranks, const uint8_t id[RTB200_UNIQUE_ID_BYTES]);

/* Sharding WITHOUT a communicator: the context renders only the bands of `rank` among `n_ranks` (same band
 * layout); nothing is exchanged between contexts.  With RTB200_F_HOST_BANDS every context writes its bands
 * into the caller's full host frame -- one process driving g GPUs from g threads needs nothing else
 * (rt_cli --gpus g); without the flag the host pointers receive the rank's compact rows.  Replaces the
 * single-process `device_ids[]` form of SURVEY.md 8b. */
int  rtb200_set_sharding(rtb200_ctx* ctx, int rank, int n_ranks);

/* The band layout rtb200_render uses with a communicator (pure host arithmetic, callable without a
 * GPU): rank `rank` renders *n_local rows and stores them compactly; out_global_rows[i] (capacity
 * >= *n_local, may be NULL) is the image row of its local row i.  band_rows 0 = default 16. */
int  rtb200_band_map(int height, int band_rows, int n_ranks, int rank,
                     int32_t* out_global_rows, int capacity, int* n_local);

/* --- helpers ------------------------------------------------------------------------------ */
/* Page-locked host memory for the caller's scene / frame buffers (so that the copies inside
 * rtb200_set_scene / rtb200_render are true async DMA). */
int  rtb200_host_alloc(size_t bytes, void** out);
void rtb200_host_free(void* p);
/* Page-locks memory the caller already owns (e.g. a shared-memory segment holding the frame of several
 * processes, RTB200_F_HOST_BANDS) for the lifetime of the registration. */
int  rtb200_host_register(void* p, size_t bytes);
int  rtb200_host_unregister(void* p);
/* FP32 FMA peak of the context's GPU measured with a register-resident FFMA chain on every SM
 * (the denominator of the roofline fraction).  Returns TFLOP/s and the SM clock it saw. */
int  rtb200_measure_fp32_peak(rtb200_ctx* ctx, double* tflops, double* sm_mhz);
/* The same measurement with its parts, so that the figure can be recomputed: detail[0] TFLOP/s,
 * [1] FFMA instructions per thread, [2] threads per block, [3] blocks, [4] elapsed ms of the best
 * launch (CUDA events), [5] median SM cycles (clock64) and [6] nanoseconds (globaltimer) one block
 * spent in the chain, [7] SM clock = cycles / ns in MHz, [8] FFMA issued per clock and SM over the whole launch
 * = [1] x [2] x [3] / ([4] x [7] x [9]) (hardware limit 128), [9] SM count.  TFLOP/s = 2 x [1] x [2] x [3] / [4].  n >= 10. */
#define RTB200_PEAK_DETAIL_COUNT 10
int  rtb200_fp32_peak_detail(rtb200_ctx* ctx, double* detail, int n);
/* Number of CUDA devices this process sees (what rt_cli --gpus g spreads its contexts over). */
int  rtb200_device_count(int* n);
/* Device properties the benchmark reports. */
int  rtb200_device_info(rtb200_ctx* ctx, int* sm_count, int* l2_bytes, char* name, int name_len);
/* Device-side stopwatch over SEVERAL calls (benchmarks time K frames as one region): records a
 * CUDA event on the context's stream in slot 0..3; elapsed = slot b - slot a in milliseconds. */
int  rtb200_timer_mark(rtb200_ctx* ctx, int slot);
int  rtb200_timer_elapsed(rtb200_ctx* ctx, int slot_a, int slot_b, double* ms);

#ifdef __cplusplus
}
#endif
#endif /* RTB200_H */
