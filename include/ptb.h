/* ptb.h - C ABI of the B200-native path-tracing hot path (libptb.so).
 *
 * Drop-in boundary for thiagoald/pathtracerpython's per-pixel, per-sample radiance estimation.
 * The reference has no FFI; its seams are plain Python calls.  Each entry point below names the
 * reference interface it replaces (file:line under the reference tree).  Host code binds these with
 * ctypes (pathtracerpython_b200/_lib.py); INTEGRATION.md shows the stub a maintainer of the
 * reference would add.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative
 * ptb_status, with text available from ptb_last_error(); no exceptions or callbacks cross the
 * boundary; pointers named d_* are DEVICE pointers owned by the caller (the Python host allocates
 * them as torch tensors), pointers named h_* / inside ptb_scene_desc are HOST pointers; every
 * launch goes to the caller's CUDA stream (a cudaStream_t passed as void*).  A scene handle is
 * immutable after creation and bound to the device that was current when it was created.
 *
 * Threads and devices: any host thread may call in, on any device and stream.  State that CUDA keeps per device
 * (kernel attributes, the SM count behind the persistent grids, the scene memory pool) is kept per device by the
 * library; per-call state (work counters, instrumentation switches) lives in the caller's workspace, so two calls
 * may run at the same time as long as they use different workspaces; ptb_last_error() is per thread.  The profiler
 * (ptb_profile_*) and the launch counter are process-wide and locked.
 */
#ifndef PTB_H
#define PTB_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PTB_OK = 0,
    PTB_ERR_ARG = -1,      /* bad argument */
    PTB_ERR_CUDA = -2,     /* CUDA runtime error (text in ptb_last_error) */
    PTB_ERR_NOMEM = -3,    /* workspace too small */
    PTB_ERR_UNSUPPORTED = -4,
    PTB_ERR_IO = -5,           /* ptb_obj_parse: file cannot be opened / read (the reference: OSError) */
    PTB_ERR_PARSE = -6,        /* ptb_obj_parse: token the reference's float()/int() rejects (ValueError) */
    PTB_ERR_PARSE_INDEX = -7,  /* ptb_obj_parse: blank last line (the reference: IndexError, scene_reader.py:31) */
    PTB_ERR_PARSE_IRREGULAR = -8 /* ptb_obj_parse: input the fast parser leaves to a line-by-line one - a vertex that does not
                                  have exactly 3 coordinates (the reference keeps any number and raises WrongDimensionError /
                                  IndexError only when a face mixes or lacks them, vector.py:63-72, 143-146), a face with fewer
                                  than 3 indices, or a face index outside the vertices read before it (Python wraps negative
                                  indices and raises IndexError beyond; scene_reader.py:81-83) */
} ptb_status;

/* Flattened scene = the data the reference keeps in `Scene` / `Obj` objects
 * (scene_reader.py:49-60, 82-88, 108-121, 171-183).  Triangles are listed in the order
 * intersect_objects visits them (main.py:91-96): every `object` in file order, the light mesh last. */
typedef struct {
    int32_t n_tri;              /* closest-hit set: object triangles + light triangles (main.py:91) */
    int32_t n_tri_obj;          /* occlusion set: object triangles only (main.py:42) */
    int32_t n_obj;              /* len(scene.objects); light triangles carry object index n_obj */
    int32_t n_light_tri;        /* == n_tri - n_tri_obj */
    const double *verts;        /* [n_tri][3][3]  Obj.triangles   (scene_reader.py:83-86) */
    const double *normals;      /* [n_tri][3]     Obj.normals     (scene_reader.py:5-8, 87) */
    const int32_t *tri_obj;     /* [n_tri]        index into scene.objects */
    const double *mat;          /* [n_obj][8]     red green blue ka kd ks kt n (scene_reader.py:175-182) */
    const double *light_areas;  /* [n_light_tri]  light_obj.areas (scene_reader.py:88) */
    double eye[3];              /* scene.eye      (scene_reader.py:152) */
    double ortho[4];            /* scene.ortho x0 y0 x1 y1 (scene_reader.py:157) */
    double ambient;             /* scene.ambient  (scene_reader.py:161) */
    double light_color[3];      /* scene.light_color[:3] (scene_reader.py:164) */
    int32_t accel;              /* PTB_ACCEL_* */
    int32_t reserved;           /* PTB_SCENE_* bits, 0 = none */
} ptb_scene_desc;

/* ptb_scene_desc.reserved: PTB_SCENE_ARRAYS_ON_DEVICE = `verts`, `normals` and `tri_obj` are DEVICE pointers (the arrays
 * reached the GPU some other way - e.g. each rank uploads 1/N of them and an all-gather over NVLink completes them,
 * dist.gather_scene_arrays); the handle copies them device-to-device.  LBVH scenes only; mat and light_areas stay
 * host pointers (they are small). */
enum { PTB_SCENE_ARRAYS_ON_DEVICE = 1 };

enum { PTB_ACCEL_AUTO = 0, PTB_ACCEL_BRUTE = 1, PTB_ACCEL_LBVH = 2 };

typedef struct ptb_scene ptb_scene;

/* Which pixels a call renders.  Ray index r = ix*H + iy (utils.py:64-69).  With d_pixels == NULL the
 * call covers the shard { r in [0, W*H) : (r / tile) % n_shards == shard } (tile-interleaved sharding
 * for multi-GPU; n_shards = 1 covers the frame).  With d_pixels != NULL it covers that explicit list
 * of n_pixels ray indices (device int32). */
typedef struct {
    int32_t width, height;      /* scene.width / scene.height (scene_reader.py:154-155) */
    int32_t shard, n_shards, tile;
    int32_t n_pixels;           /* only with d_pixels */
    const int32_t *d_pixels;
} ptb_pixels;

enum { PTB_PREC_FP32 = 0, PTB_PREC_FP64 = 1 };
enum {
    PTB_FLAG_NO_FP64_RETRACE = 1,   /* FP32 only: skip the FP64 re-trace of uncertain primary hits */
    PTB_FLAG_SEPARATE_KERNELS = 2,  /* A/B and test switch: one primary ray per SAMPLE (default: once per pixel, they are
                                       identical, main.py:191) and, for small scenes, the separate shade / shadow /
                                       closest-hit kernels instead of the fused per-bounce kernel */
    PTB_FLAG_COUNT_NODES = 4,       /* instrumented run: count LBVH nodes visited / triangles tested */
    PTB_FLAG_PHYSICAL = 8,          /* physical mode (not in the reference): ray queries (t > 0), two-sided
                                       normals with a true tangent frame, kd / ks / kt as diffuse / Phong / refractive lobes,
                                       area-uniform NEE with the geometry term, RGB throughput, Russian roulette */
    PTB_FLAG_NO_ROULETTE = 16       /* physical mode: fixed depth, no Russian roulette (unbiasedness check) */
};

typedef struct {
    ptb_pixels px;
    int32_t sample_begin, sample_end;   /* samples [s0, s1)  = main.py:186 `rays_counter`, -r */
    int32_t bounces;                    /* main.py:192, -b */
    uint32_t seed;                      /* Philox key (the reference's RNG is unseeded, A-Q18) */
    int32_t precision;                  /* PTB_PREC_* */
    int32_t flags;                      /* PTB_FLAG_* */
    int32_t samples_per_batch;          /* 0 = choose from the workspace size */
    int32_t reserved;
} ptb_render_params;

/* Optional per-(pixel, sample, bounce) records for parity check 2; all device pointers, any may be
 * NULL.  Index = ((pixel_slot * n_samples) + (sample - sample_begin)) * bounces + bounce, where
 * pixel_slot is the position of the pixel in the call's pixel set. */
typedef struct {
    int32_t *d_tri;        /* closest-hit triangle id, -1 miss, -2 path already ended */
    float *d_contrib;      /* [..][3] new_color * accumulated_k (main.py:230-231) */
    float *d_k_after;      /* accumulated_k after the bounce (main.py:251, 263) */
    int8_t *d_lobe;        /* 0 diffuse, 1 specular, -1 none (main.py:241) */
    int8_t *d_occ;         /* bit s set = light sample s occluded (main.py:49-51) */
    int8_t *d_leak;        /* object index the NEE colour is taken from (main.py:42, 70) */
    double *d_contrib64;   /* FP64 mode: same as d_contrib in double */
    double *d_k_after64;
} ptb_path_records;

typedef struct {
    uint64_t closest_rays;   /* closest-hit rays traced (intersect_objects calls on live paths) */
    uint64_t shadow_rays;    /* shadow rays traced (3 per compute_color call) */
    uint64_t shaded;         /* compute_color calls */
    uint64_t samples;        /* pixel samples started */
    uint64_t retraced;       /* primary rays re-traced in FP64 */
    uint64_t launches;       /* kernels launched by this library since ptb_counters_reset */
    uint64_t nodes_visited;  /* PTB_FLAG_COUNT_NODES runs only: LBVH nodes fetched by closest-hit rays */
    uint64_t tris_tested;    /* ... triangles tested by closest-hit rays */
    uint64_t nodes_visited_shadow;
    uint64_t tris_tested_shadow;
} ptb_counters;

const char *ptb_last_error(void);
int ptb_version(void);

/* replaces Scene(path) object hand-off into main() (main.py:167) + the per-task pickling of the
 * scene (main.py:199-201): uploads the flattened scene and builds the acceleration structure. */
int ptb_scene_create(const ptb_scene_desc *desc, void *stream, ptb_scene **out);
int ptb_scene_destroy(ptb_scene *scene);
/* ptb_scene_destroy waits for the streams the handle was used on (not for the device) and frees in stream order.
 * Scene memory is taken from a stream-ordered pool PRIVATE to the library (one per device; the device's default pool is
 * left alone) and returns to it on ptb_scene_destroy, so that creating a scene per frame does not pay the driver's
 * allocator each time.  The pool keeps at most 16 GiB of unused memory across synchronisations; ptb_pool_trim hands
 * all of it back to the driver (synchronises the device). */
int ptb_pool_trim(void);
int ptb_scene_info(const ptb_scene *scene, int32_t *accel, int32_t *n_nodes, double *build_ms);
/* 1 when the scene's hierarchies carry supporting-plane ids (an LBVH scene with at most 255 distinct planes) and the
 * traversal kernels skip subtrees lying in the plane a ray starts on; 0 otherwise (brute-force scenes, soups). */
int ptb_scene_plane_culling(const ptb_scene *scene);

/* Replicating a built scene to other GPUs instead of rebuilding it (the reference pickles the whole scene into
 * every pool task, main.py:199-201, 208-216; here one rank builds and the others receive).
 * ptb_bvh_export copies both hierarchies of an LBVH scene (closest-hit tree over object + light triangles and the
 * occlusion tree over object triangles: 64 B nodes, leaf-ordered 64 B triangle records, slot -> reference index; plus
 * the per-triangle supporting-plane ids the nodes refer to)
 * into ONE caller-owned device buffer of ptb_bvh_export_bytes(scene) bytes (0 for a scene without LBVH) that can
 * travel through a single broadcast.  ptb_bvh_import is ptb_scene_create with the sort / tree / refit kernels
 * replaced by a device-to-device copy of such a buffer: `desc` must describe the same scene (checked against the
 * buffer's header: triangle counts, box padding, format version), desc->accel is taken as PTB_ACCEL_LBVH.  The
 * build is deterministic, so an imported scene renders bit-identically to a built one. */
size_t ptb_bvh_export_bytes(const ptb_scene *scene);
int ptb_bvh_export(const ptb_scene *scene, void *d_blob, size_t blob_bytes, void *stream);
int ptb_bvh_import(const ptb_scene_desc *desc, const void *d_blob, size_t blob_bytes, void *stream, ptb_scene **out);

/* replaces make_screen_pts + make_rays + one round of intersect_objects over all primary rays
 * (utils.py:55-69, main.py:191, 199-204, 83-122).  d_tri[i]: global triangle id or -1; d_t[i]:
 * signed ray parameter of the hit (unit direction) or 0.  n outputs = size of the pixel set. */
size_t ptb_trace_primary_workspace(int64_t n_pixels, int32_t precision);   /* bytes ptb_trace_primary needs */
int ptb_trace_primary(const ptb_scene *scene, const ptb_pixels *px, int32_t precision, int32_t flags,
                      int32_t *d_tri, float *d_t, void *d_workspace, size_t workspace_bytes, void *stream);

/* closest hit for caller-supplied rays (intersect_objects, main.py:83-122): d_rays [n][6] (origin,
 * direction; float for FP32, double for FP64). */
int ptb_trace_closest(const ptb_scene *scene, const void *d_rays, int64_t n, int32_t precision,
                      int32_t *d_tri, void *d_t, void *stream);

/* ptb_trace_closest for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * arguments, same results as ptb_trace_closest (bit for bit); the two ray queues and the count word that
 * ptb_trace_closest takes from cudaMallocAsync live in a file-scope `static __device__` array of the library instead
 * (2 x 32 B x PTB_TRACE_CLOSEST_NOALLOC_MAX_RAYS + 256 B = 8 MiB + 256 B, held by every GPU that loads the library
 * until the process exits).  The call makes no cudaMalloc / cudaFree / cudaMallocAsync and calls no allocating library
 * routine; more than PTB_TRACE_CLOSEST_NOALLOC_MAX_RAYS rays in one call is refused with PTB_ERR_NOMEM and a message
 * (split the call).  Every call on a device shares the one array, which keeps the last call's contents (the call
 * clears the word that must start at zero): calls that overlap on the GPU - two streams, two host threads - RACE;
 * keep them on one stream or order them with events.  LBVH scenes are walked by the static-assignment kernel (the
 * dynamic-fetch kernel needs a work counter from a workspace). */
#define PTB_TRACE_CLOSEST_NOALLOC_MAX_RAYS 131072
int ptb_trace_closest_noalloc(const ptb_scene *scene, const void *d_rays, int64_t n, int32_t precision,
                              int32_t *d_tri, void *d_t, void *stream);

/* replaces the sample loop x bounce loop of main() (main.py:186-271): for every pixel of the set and
 * every sample in [s0, s1) it ADDS the sample's radiance to d_accum[pixel_slot] (float4: r, g, b,
 * sample count).  The mean (main.py:274-280) is taken by ptb_tonemap / the caller. */
size_t ptb_trace_paths_workspace(const ptb_render_params *p, int32_t samples_per_batch);
int ptb_trace_paths(const ptb_scene *scene, const ptb_render_params *p, void *d_workspace,
                    size_t workspace_bytes, float *d_accum /* [n_pixel_slots][4] */,
                    const ptb_path_records *records /* may be NULL */, void *stream);

/* ptb_trace_paths with the sample index shifted by a DEVICE word: sample s of the call is sample s + *d_sample_offset of the
 * stream (main.py:186 `rays_counter`; no counterpart in the reference, whose sample loop lives on the host).  The launch
 * sequence of a call makes no allocation, no copy from host memory and no synchronisation, so it can be recorded into a CUDA
 * graph (stream capture); with the offset in device memory - and advanced inside the same graph - every replay adds the
 * NEXT samples of every pixel to d_accum: progressive rendering of a launch-bound frame (the shipped 40 x 40 scene: fifteen
 * launches per frame) at the cost of one graph launch.  pathtracerpython_b200.render.ProgressiveFrame is that graph. */
int ptb_trace_paths_at(const ptb_scene *scene, const ptb_render_params *p, void *d_workspace, size_t workspace_bytes,
                       float *d_accum, const ptb_path_records *records, const int32_t *d_sample_offset, void *stream);

/* replaces make_image (utils.py:150-161): d_accum is the FULL frame in ray-index order
 * (slot = r); writes the H x W x 3 u8 image, rows top to bottom. d_workspace >= 64 bytes. */
int ptb_tonemap(const float *d_accum, int32_t width, int32_t height, uint8_t *d_rgb,
                void *d_workspace, void *stream);

/* Opt-in display transform (no counterpart in the reference's code path: it parses the scene keywords `background`
 * and `tonemapping`, scene_reader.py:158-159, 167-168, and never reads them; ptb_tonemap above is what the reference
 * does).  Per channel: mean / (mean + tonemapping) (tonemapping == 0: the mean itself), clamped to [0, 1], * 255
 * truncated; pixels with d_primary_tri[r] < 0 (primary ray left the scene; d_primary_tri may be NULL) show
 * `background` (3 doubles in [0, 1], NULL = black).  Same frame layout as ptb_tonemap. */
int ptb_tonemap_scene(const float *d_accum, const int32_t *d_primary_tri, int32_t width, int32_t height,
                      double tonemapping, const double *background, uint8_t *d_rgb, void *stream);

/* scatter a shard's accumulation buffer (pixel-slot order) into a full-frame buffer (ray-index order) */
int ptb_scatter_accum(const ptb_pixels *px, const float *d_accum_shard, float *d_accum_full, void *stream);

/* Per-kernel-class device timing (CUDA events on the launch stream), for the roofline report.
 * kinds: 0 camera rays, 1 closest-hit, 2 FP64 re-trace, 3 shade/bounce, 4 shadow/any-hit + NEE,
 * 5 batch reduce, 6 tone map, 7 other, 8 LBVH build, 9 per-bounce ray sort,
 * 10 fused per-bounce kernel (shade + shadow + next closest hit, small scenes).  ptb_profile_read synchronises the device,
 * writes the accumulated milliseconds and launch counts per kind (arrays of PTB_PROF_KINDS) and
 * clears the accumulators. */
#define PTB_PROF_KINDS 11
int ptb_profile_enable(int on);
int ptb_profile_read(double *ms_per_kind, uint64_t *launches_per_kind);

/* FP32 FMA issue peak of the current device, measured with a register-resident FMA microbenchmark
 * (roofline denominator for the compute-bound small-scene kernels). */
int ptb_fp32_peak(double *tflops, void *stream);

/* replaces Obj(path) (scene_reader.py:49-104; line cleaning :29-46) for meshes of 10^6..10^7 triangles:
 * multi-threaded host parse of `v` / `f` lines with the reference's rules (comments, tabs, polygons fanned
 * about their first vertex, negative = relative indices, 'i/j/k' rejected).  On PTB_OK *verts holds
 * 3*n_verts doubles (bit-identical to Python's float()), *faces 3*n_faces vertex indices, *skipped (may be
 * passed as NULL) the cleaned text of every line with an unknown keyword, '\n'-separated, for the caller to
 * report as the reference does (:102-104).  Irregular input (PTB_ERR_PARSE_IRREGULAR) and malformed tokens are reported,
 * not repaired: pathtracerpython_b200.scene.Obj then re-reads the file line by line and raises what the reference
 * raises for the first offending line (tests/golden/obj_corners.json).  All three are released with ptb_obj_free.  threads <= 0: all
 * host cores.  Pure host code: needs no device. */
int ptb_obj_parse(const char *path, double **verts, int64_t *n_verts, int64_t **faces, int64_t *n_faces,
                  char **skipped, int threads);
void ptb_obj_free(void *p);
/* Companion writer for generated meshes (no reference counterpart: the reference only reads): `v` lines with the
 * shortest decimals that parse back to the same doubles, then 1-based `f` lines.  header (may be NULL) becomes a
 * leading `# ...` comment. */
int ptb_obj_write(const char *path, const double *verts, int64_t n_verts, const int64_t *faces, int64_t n_faces,
                  const char *header);

/* 64 B gather peak of the current device (roofline denominator of the LBVH kernels, which are bound by scattered node
 * and triangle-record fetches, not by DRAM bytes): every lane of a traversal-shaped launch (128 threads, 10 blocks per
 * SM) chases its own chain of 64 B records - two 256-bit loads, the next index taken from the record just read -
 * through a buffer of `buffer_bytes`.  64 KB measures L1, 64 MB L2, > 1 GB HBM.  Result in 10^9 fetches per second. */
int ptb_gather_peak(size_t buffer_bytes, double *gfetch_per_s, void *stream);
/* The same chase with the records staged in SHARED memory (bytes_per_block of them in each of the 10 blocks per SM;
 * 1 KB .. 20 KB): the rate a walk would get from tree levels staged there ("why the top of the tree is not
 * staged"). */
int ptb_gather_peak_smem(size_t bytes_per_block, double *gfetch_per_s, void *stream);

/* Known-answer entry for the device functions that have no seam of their own (tier 1).  d_in / d_out are
 * DEVICE arrays of doubles, n items:
 *   what 0  rotate about (0,1,0) (main.py:148-162):           in [n][4] = angle, v.xyz           -> out [n][3]
 *   what 1  pick_random_triangle + sample_random_pt (utils.py:21-46) under the injected Philox stream:
 *                                                             in [n][4] = pixel, sample, bounce, quad -> out [n][4] = pick, point.xyz
 *   what 2  make_screen_pts + make_rays (utils.py:55-69):     in [n] = ray index (width, height) -> out [n][3] direction as returned
 * precision PTB_PREC_FP64: the reference-order flavour; PTB_PREC_FP32: the product's device functions. */
int ptb_kat(const ptb_scene *scene, int32_t what, int32_t precision, const double *d_in, int64_t n, double *d_out,
            int32_t width, int32_t height, uint32_t seed, void *stream);

int ptb_counters_read(const ptb_scene *scene, ptb_counters *out, void *stream); /* synchronises stream */
int ptb_counters_reset(const ptb_scene *scene, void *stream);
int64_t ptb_pixel_count(const ptb_pixels *px);

#ifdef __cplusplus
}
#endif
#endif /* PTB_H */
