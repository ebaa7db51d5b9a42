// ptb_api.cu - C ABI (include/ptb.h) over the wavefront kernels.  Host logic only: scene upload and
// precompute, workspace carving, the per-batch / per-bounce launch sequence.
#include <math.h>
#include <stdarg.h>

#include <algorithm>
#include <chrono>
#include <vector>

#include "ptb_image.cuh"
#include "ptb_shade64.cuh"
#include "ptb_planes.cuh"
#include "ptb_bvh_build.cuh"
#include "ptb_bvh_kernels.cuh"
#include "ptb_bvh_dyn.cuh"
#include "ptb_fused.cuh"
#include "ptb_fused2.cuh"
#include "ptb_phys.cuh"

namespace ptb {

static thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};
std::mutex g_state_mu;
static DeviceState g_dev[PTB_MAX_DEVICES];

DeviceState *device_state() {
    int dev = 0;
    cudaGetDevice(&dev);
    DeviceState *ds = &g_dev[(unsigned)dev < PTB_MAX_DEVICES ? dev : 0];
    if (!ds->ready.load(std::memory_order_acquire)) {
        std::lock_guard<std::mutex> lk(g_state_mu);
        if (!ds->ready.load(std::memory_order_relaxed)) {
            int sms = 0;
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            ds->sms = sms > 0 ? sms : 148;
            // Scene memory comes from a PRIVATE stream-ordered pool: creating and destroying scenes (a scene per frame,
            // bench.py's end-to-end step) recycles pool memory instead of going to the driver's allocator (cudaFree
            // synchronises the device and was measured at 4-240 ms per scene), and the device's default pool - which other
            // cudaMallocAsync users of the process share - keeps its own release threshold.  The pool holds on to at most
            // PTB_POOL_KEEP_BYTES of unused memory across synchronisations (two 10 M-triangle scenes with a build arena are
            // 9 GB); ptb_pool_trim hands everything back.
            ds->pool = nullptr;
            cudaMemPoolProps props;
            memset(&props, 0, sizeof(props));
            props.allocType = cudaMemAllocationTypePinned;
            props.handleTypes = cudaMemHandleTypeNone;
            props.location.type = cudaMemLocationTypeDevice;
            props.location.id = dev;
            cudaMemPool_t pool = nullptr;
            if (cudaMemPoolCreate(&pool, &props) == cudaSuccess) {
                unsigned long long keep = PTB_POOL_KEEP_BYTES;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
                ds->pool = pool;
            } else {
                cudaGetLastError();
            }
            // PTB_L2_PERSIST_MB=n: set n MB of L2 aside for persisting lines; the LBVH kernels then pin their node array there
            // (l2_window, ptb_api_render.inl).  Off by default: measured, see DESIGN.md section 4.
            ds->l2_persist_bytes = 0; ds->l2_window_max = 0;
            if (const char *mb = getenv("PTB_L2_PERSIST_MB")) {
                cudaDeviceProp prop;
                if (atoi(mb) > 0 && cudaGetDeviceProperties(&prop, dev) == cudaSuccess && prop.persistingL2CacheMaxSize > 0) {
                    const size_t want = std::min<size_t>((size_t)atoi(mb) << 20, (size_t)prop.persistingL2CacheMaxSize);
                    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
                        ds->l2_persist_bytes = want; ds->l2_window_max = (size_t)prop.accessPolicyMaxWindowSize;
                    } else cudaGetLastError();
                }
            }
            ds->ready.store(true, std::memory_order_release);
        }
    }
    return ds;
}

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static int num_sms() { return device_state()->sms; }
// persistent grid: a multiple of the SM count, capped by the work available
static int grid_for(size_t work, int threads, int blocks_per_sm) {
    size_t need = (work + threads - 1) / threads;
    size_t cap = (size_t)num_sms() * blocks_per_sm;
    return (int)std::max<size_t>(1, std::min(need, cap));
}

// ---- optional per-kernel-class timing with CUDA events on the launch stream
enum { PK_GEN = 0, PK_CLOSEST, PK_RETRACE, PK_SHADE, PK_SHADOW, PK_REDUCE, PK_TONEMAP, PK_OTHER, PK_BUILD, PK_SORT, PK_BOUNCE };
struct ProfRec { int kind; cudaEvent_t a, b; };
static std::atomic<bool> g_prof_on{false};
static std::vector<ProfRec> g_prof;             // guarded by g_state_mu (any host thread may be launching)
static std::vector<cudaEvent_t> g_event_pool;   // guarded by g_state_mu
static cudaEvent_t prof_event() {
    cudaEvent_t e;
    {
        std::lock_guard<std::mutex> lk(g_state_mu);
        if (!g_event_pool.empty()) { e = g_event_pool.back(); g_event_pool.pop_back(); return e; }
    }
    cudaEventCreate(&e);
    return e;
}
struct ProfScope {
    bool on; ProfRec r; cudaStream_t st;
    ProfScope(int kind, cudaStream_t s) : on(g_prof_on.load(std::memory_order_relaxed)), st(s) {
        if (on) { r.kind = kind; r.a = prof_event(); r.b = prof_event(); cudaEventRecord(r.a, st); }
    }
    ~ProfScope() { if (on) { cudaEventRecord(r.b, st); std::lock_guard<std::mutex> lk(g_state_mu); g_prof.push_back(r); } }
};

}  // namespace ptb

using namespace ptb;

extern "C" int ptb_profile_enable(int on) { g_prof_on.store(on != 0); return PTB_OK; }
extern "C" int ptb_profile_read(double *ms, uint64_t *launches) {
    PTB_CUDA(cudaDeviceSynchronize());
    for (int i = 0; i < PTB_PROF_KINDS; ++i) { if (ms) ms[i] = 0.0; if (launches) launches[i] = 0; }
    std::lock_guard<std::mutex> lk(g_state_mu);
    for (ProfRec &r : g_prof) {
        float t = 0.f;
        cudaEventElapsedTime(&t, r.a, r.b);
        if (ms) ms[r.kind] += t;
        if (launches) launches[r.kind] += 1;
        g_event_pool.push_back(r.a); g_event_pool.push_back(r.b);
    }
    g_prof.clear();
    return PTB_OK;
}

// Scene memory: the device's private pool (device_state()), plain cudaMallocAsync if the pool could not be created
static cudaError_t scene_malloc(void **p, size_t bytes, cudaStream_t st) {
    DeviceState *ds = device_state();
    if (ds->pool) return cudaMallocFromPoolAsync(p, std::max<size_t>(bytes, 256), ds->pool, st);
    return cudaMallocAsync(p, std::max<size_t>(bytes, 256), st);
}
// ptb_pool_trim (include/ptb.h): hand the pool's unused memory back to the driver
extern "C" int ptb_pool_trim(void) {
    DeviceState *ds = device_state();
    PTB_CUDA(cudaDeviceSynchronize());
    if (ds->pool) PTB_CUDA(cudaMemPoolTrimTo(ds->pool, 0));
    return PTB_OK;
}

struct ptb_scene {
    DScene d;                       // device view handed to kernels by value
    std::vector<void *> allocs;     // device allocations owned by the handle
    mutable std::mutex mu;          // guards `streams`
    mutable std::vector<cudaStream_t> streams;   // every stream a call has used the scene on (ptb_scene_destroy waits for them)
    cudaStream_t create_stream;
    int device;
    float dt_scale;                 // FP32 error scale of the ray parameter (see ptb_isect.cuh)
    double build_ms;
    int n_nodes;
    PlaneScene planes;              // plane-grouped records (brute-force scenes)
    float pad;                      // LBVH box inflation
    bool use_planes;
    bool fused_x2;                  // packed-FP32 flavour of the fused per-bounce kernel (ptb_fused2.cuh); PTB_FUSED_X2=0: scalar
    bool bvh_dynamic;               // persistent kernels with dynamic ray fetch (ptb_bvh_dyn.cuh)
    bool sort_rays;                 // per-bounce Morton re-ordering of the live paths (ptb_sort.cuh)
};

template <typename T> static int upload(ptb_scene *s, const std::vector<T> &h, const T **out, cudaStream_t st) {
    void *p = nullptr;
    size_t bytes = std::max<size_t>(h.size(), 1) * sizeof(T);
    PTB_CUDA(scene_malloc(&p, bytes, st));
    s->allocs.push_back(p);
    if (!h.empty()) PTB_CUDA(cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, st));
    *out = (const T *)p;
    return PTB_OK;
}

template <typename T> static int upload_raw(ptb_scene *s, const T *src, size_t count, const T **out, cudaStream_t st,
                                            cudaMemcpyKind kind = cudaMemcpyHostToDevice) {
    void *p = nullptr;
    PTB_CUDA(scene_malloc(&p, std::max<size_t>(count, 1) * sizeof(T), st));
    s->allocs.push_back(p);
    if (count) PTB_CUDA(cudaMemcpyAsync(p, src, count * sizeof(T), kind, st));
    *out = (const T *)p;
    return PTB_OK;
}

static inline float bits_f(int v) { float f; memcpy(&f, &v, 4); return f; }
static inline void cross_h(const double *a, const double *b, double *c) {
    c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0];
}
static inline double dot_h(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// FP32 traversal record of one triangle (layout: ptb_scene.cuh)
static TriRec make_trirec(const double *v, int obj, double scale) {
    const double *A = v, *B = v + 3, *C = v + 6;
    double ab[3], cb[3], n[3], e1[3], e2[3], N[3], mu[3], mv[3];
    for (int i = 0; i < 3; ++i) { ab[i] = A[i] - B[i]; cb[i] = C[i] - B[i]; e1[i] = B[i] - A[i]; e2[i] = C[i] - A[i]; }
    cross_h(ab, cb, n);                                  // utils.py:112-113
    double nl = sqrt(dot_h(n, n));
    for (int i = 0; i < 3; ++i) n[i] /= nl;
    cross_h(e1, e2, N);
    double inv = 1.0 / dot_h(N, N);
    cross_h(e2, N, mu);
    cross_h(N, e1, mv);
    for (int i = 0; i < 3; ++i) { mu[i] *= inv; mv[i] *= inv; }
    double mw[3] = {mu[0] + mv[0], mu[1] + mv[1], mu[2] + mv[2]};
    double g = std::max(sqrt(dot_h(mu, mu)), std::max(sqrt(dot_h(mv, mv)), sqrt(dot_h(mw, mw))));
    TriRec r;
    r.q0 = make_float4((float)n[0], (float)n[1], (float)n[2], (float)dot_h(n, A));
    r.q1 = make_float4((float)mu[0], (float)mu[1], (float)mu[2], (float)(-dot_h(A, mu)));
    r.q2 = make_float4((float)mv[0], (float)mv[1], (float)mv[2], (float)(-dot_h(A, mv)));
    r.q3 = make_float4((float)(32.0 * 5.9604644775390625e-08 * scale * g), 0.f, bits_f(obj), 0.f);
    return r;
}

// Group the triangles of a small scene by supporting plane (ptb_planes.cuh).  Groups are ordered by
// their lowest reference index, triangles inside a group keep reference order (object triangles before
// light triangles because the light mesh is last).
static void build_plane_groups(const ptb_scene_desc *desc, const std::vector<TriRec> &tri, std::vector<PlaneRec> &planes,
                               int &n_occ) {
    const int n = desc->n_tri;
    struct G { double n[3], d; std::vector<int> tris; };
    std::vector<G> groups;
    double scale = 1.0;
    for (size_t i = 0; i < (size_t)n * 9; ++i) scale = std::max(scale, fabs(desc->verts[i]));
    for (int i = 0; i < n; ++i) {
        const double *A = desc->verts + 9 * (size_t)i, *B = A + 3, *C = A + 6;
        double ab[3], cb[3], nn[3];
        for (int k = 0; k < 3; ++k) { ab[k] = A[k] - B[k]; cb[k] = C[k] - B[k]; }
        cross_h(ab, cb, nn);
        double nl = sqrt(dot_h(nn, nn));
        int big = 0;
        for (int k = 1; k < 3; ++k) if (fabs(nn[k]) > fabs(nn[big])) big = k;
        double sgn = nn[big] < 0 ? -1.0 : 1.0;
        for (int k = 0; k < 3; ++k) nn[k] = sgn * nn[k] / nl;
        double d = dot_h(nn, A);
        int found = -1;
        // pair only triangles of the same set (object / light) so a group is wholly in or out of the occlusion set
        if (nl > 0 && nl == nl)
            for (size_t g = 0; g < groups.size(); ++g) {
                const G &q = groups[g];
                if (q.tris.size() < 2 && ((q.tris[0] < desc->n_tri_obj) == (i < desc->n_tri_obj)) &&
                    fabs(q.n[0] - nn[0]) < 1e-10 && fabs(q.n[1] - nn[1]) < 1e-10 && fabs(q.n[2] - nn[2]) < 1e-10 &&
                    fabs(q.d - d) < 1e-10 * scale) { found = (int)g; break; }
            }
        if (found < 0) { G q; q.n[0] = nn[0]; q.n[1] = nn[1]; q.n[2] = nn[2]; q.d = d; groups.push_back(q); found = (int)groups.size() - 1; }
        groups[found].tris.push_back(i);
    }
    // groups are created in order of their first triangle => object groups precede light groups
    n_occ = 0;
    for (const G &q : groups) {
        PlaneRec pr;
        const int ta = q.tris[0], tb = q.tris.size() > 1 ? q.tris[1] : -1;
        // the plane stored is triangle A's own (FP32-rounded) plane so a single-triangle group is exactly TriRec
        pr.nd = tri[ta].q0;
        pr.a1 = tri[ta].q1; pr.a2 = tri[ta].q2;
        if (tb >= 0) { pr.b1 = tri[tb].q1; pr.b2 = tri[tb].q2; }
        else { pr.b1 = make_float4(0.f, 0.f, 0.f, -1.f); pr.b2 = make_float4(0.f, 0.f, 0.f, -1.f); }
        pr.ref_a = ta; pr.ref_b = tb; pr.kind = 0;
        if (tb >= 0) {   // parallelogram: B's vertex set == {a1, a3, a1 + a3 - a2}
            const double *A = desc->verts + 9 * (size_t)ta, *B = desc->verts + 9 * (size_t)tb;
            double want[3][3];
            for (int k = 0; k < 3; ++k) { want[0][k] = A[k]; want[1][k] = A[6 + k]; want[2][k] = A[k] + A[6 + k] - A[3 + k]; }
            bool used[3] = {false, false, false};
            int matched = 0;
            for (int w = 0; w < 3; ++w)
                for (int v = 0; v < 3; ++v) {
                    if (used[v]) continue;
                    const double tol = (w < 2 ? 0.0 : 1e-12 * scale);
                    if (fabs(B[3 * v] - want[w][0]) <= tol && fabs(B[3 * v + 1] - want[w][1]) <= tol && fabs(B[3 * v + 2] - want[w][2]) <= tol) {
                        used[v] = true; ++matched; break;
                    }
                }
            if (matched == 3) pr.kind = 1;
        }
        pr.n_obj = ta < desc->n_tri_obj ? (tb >= 0 ? 2 : 1) : 0;
        if (pr.n_obj) n_occ++;
        planes.push_back(pr);
    }
}

// GPU LBVH build (ptb_bvh_build.cuh).  Temporary buffers are freed before returning; the handle keeps
// the 64 B nodes, the leaf-ordered TriRec array and the slot -> reference index table.
// Builds over the first `n` triangles of the scene (reference order): n = n_tri gives the closest-hit tree,
// n = n_tri_obj the occlusion tree (main.py:42 loops over scene.objects only; the light mesh is last).
// The hierarchy is object-major: triangles are ordered by (rank of their object, Morton code inside the object's own
// box) and the radix tree is built over that 64-bit key, so the top levels separate whole meshes (the reference's
// scenes are lists of `object file.obj` meshes: walls, blocks) and every mesh gets a subtree with tight boxes - flat
// for a wall.  n_objs = number of object ids among the n triangles.
static size_t lbvh_arena_bytes(int n, int n_objs) {      // upper bound of what build_lbvh carves (sum of its temporaries)
    const size_t nn = (size_t)n, tiles = (nn + PS_TILE - 1) / PS_TILE;
    return nn * (2 * sizeof(Aabb) + sizeof(float4) + sizeof(int2) + 4 * 9 + 8 + 1) + 4 * 256 * tiles + (size_t)n_objs * 52 + 64 * 256;
}

static int build_lbvh(ptb_scene *s, int n, int n_objs, const TriRec *src_tri, const BvhNode **nodes_out, const TriRec **tri_out,
                      const int **ids_out, double *ms_out, void *arena, size_t arena_bytes, cudaStream_t st) {
    DScene &d = s->d;
    const int n_tiles = (n + PS_TILE - 1) / PS_TILE;
    Aabb *tri_box = nullptr, *node_box = nullptr; float4 *centroid = nullptr; unsigned *scene_box = nullptr, *keys[2] = {nullptr, nullptr};
    int *vals[2] = {nullptr, nullptr}, *parent_inner = nullptr, *parent_leaf = nullptr, *node_minref = nullptr; int2 *child = nullptr;
    unsigned *ghist = nullptr, *flags = nullptr, *totals = nullptr, *d_n = nullptr; BvhNode *nodes = nullptr; TriRec *tri_sorted = nullptr;
    unsigned *obj_box = nullptr, *morton_orig = nullptr; unsigned long long *key64 = nullptr; float *obj_lo = nullptr, *obj_inv = nullptr; int *rank = nullptr;
    unsigned char *node_pid = nullptr;
    // temporaries are carved out of ONE arena owned by the caller (both trees of a scene share it): twenty-odd cudaMalloc /
    // cudaFree pairs per tree cost ten times the 1.6 ms the build kernels take (cudaFree synchronises the device)
    struct Want { void **p; size_t bytes; bool keep; };
    const Want wants[] = {
        {(void **)&tri_box, sizeof(Aabb) * (size_t)n, false}, {(void **)&node_box, sizeof(Aabb) * (size_t)n, false},
        {(void **)&centroid, sizeof(float4) * (size_t)n, false}, {(void **)&scene_box, 64, false},
        {(void **)&keys[0], 4 * (size_t)n, false}, {(void **)&keys[1], 4 * (size_t)n, false},
        {(void **)&vals[0], 4 * (size_t)n, false}, {(void **)&vals[1], 4 * (size_t)n, true},
        {(void **)&parent_inner, 4 * (size_t)n, false}, {(void **)&parent_leaf, 4 * (size_t)n, false},
        {(void **)&node_minref, 4 * (size_t)n, false}, {(void **)&child, sizeof(int2) * (size_t)n, false},
        {(void **)&ghist, 4 * (size_t)256 * n_tiles, false}, {(void **)&totals, 4 * 256, false},
        {(void **)&d_n, 16, false}, {(void **)&flags, 4 * (size_t)n, false},
        {(void **)&obj_box, 24 * (size_t)n_objs, false}, {(void **)&morton_orig, 4 * (size_t)n, false},
        {(void **)&key64, 8 * (size_t)n, false}, {(void **)&obj_lo, 12 * (size_t)n_objs, false},
        {(void **)&obj_inv, 12 * (size_t)n_objs, false}, {(void **)&rank, 4 * (size_t)n_objs, false},
        {(void **)&node_pid, (size_t)n, false},
        {(void **)&nodes, sizeof(BvhNode) * (size_t)n, true}, {(void **)&tri_sorted, sizeof(TriRec) * (size_t)n, true}};
    size_t at = 0;
    for (const Want &w : wants) {
        if (w.keep) {
            PTB_CUDA(scene_malloc(w.p, w.bytes, st));
            s->allocs.push_back(*w.p);
        } else {
            *w.p = (char *)arena + at;
            at += (std::max<size_t>(w.bytes, 16) + 255) & ~(size_t)255;
        }
    }
    if (at > arena_bytes) { set_error("LBVH build: arena of %zu bytes, %zu needed", arena_bytes, at); return PTB_ERR_NOMEM; }
    struct Events {         // destroyed on every way out
        cudaEvent_t a, b;
        Events() { cudaEventCreate(&a); cudaEventCreate(&b); }
        ~Events() { cudaEventDestroy(a); cudaEventDestroy(b); }
    } ev;
    const cudaEvent_t e0 = ev.a, e1 = ev.b;
    const unsigned h_n = (unsigned)n;
    cudaMemcpyAsync(d_n, &h_n, sizeof(unsigned), cudaMemcpyHostToDevice, st);
    cudaEventRecord(e0, st);
    {
        ProfScope ps_(PK_BUILD, st);
        const int g = grid_for((size_t)n, 256, 8);
        k_bounds_init<<<1, 32, 0, st>>>(scene_box);
        k_obj_bounds_init<<<grid_for((size_t)n_objs * 6, 256, 1), 256, 0, st>>>(obj_box, n_objs);
        k_tri_bounds<<<g, 256, 0, st>>>(d.verts64, d.tri_obj, n, tri_box, centroid, scene_box, obj_box);
        g_launches += 3;
        // object boxes -> host: Morton normalisation per object and the rank of every object along the Morton curve of
        // the object centres (a spatial order for the top levels); PTB_BVH_FLAT=1 keeps one scene-wide Morton order
        std::vector<unsigned> h_box((size_t)n_objs * 6 + 6);
        cudaMemcpyAsync(h_box.data(), obj_box, 24 * (size_t)n_objs, cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(h_box.data() + (size_t)n_objs * 6, scene_box, 24, cudaMemcpyDeviceToHost, st);
        if (cudaStreamSynchronize(st) != cudaSuccess) { set_error("LBVH build: bounds pass failed"); return PTB_ERR_CUDA; }
        auto unorder = [](unsigned o) { unsigned b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o; float f; memcpy(&f, &b, 4); return f; };
        const bool flat = getenv("PTB_BVH_FLAT") != nullptr;
        float slo[3], shi[3];
        for (int k = 0; k < 3; ++k) { slo[k] = unorder(h_box[(size_t)n_objs * 6 + k]); shi[k] = unorder(h_box[(size_t)n_objs * 6 + 3 + k]); }
        std::vector<float> h_lo((size_t)n_objs * 3), h_inv((size_t)n_objs * 3);
        std::vector<int> h_rank(n_objs, 0);
        std::vector<std::pair<unsigned, int>> order;
        auto ex10 = [](unsigned v) { v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu; v = (v * 0x00000011u) & 0xC30C30C3u; return (v * 0x00000005u) & 0x49249249u; };
        for (int o = 0; o < n_objs; ++o) {
            unsigned code = 0;
            float ext = 1e-30f, blo3[3], bhi3[3];
            for (int k = 0; k < 3; ++k) {
                blo3[k] = unorder(h_box[(size_t)o * 6 + k]); bhi3[k] = unorder(h_box[(size_t)o * 6 + 3 + k]);
                if (flat || !(bhi3[k] >= blo3[k])) { blo3[k] = slo[k]; bhi3[k] = shi[k]; }   // empty object: any box will do
                ext = std::max(ext, bhi3[k] - blo3[k]);
            }
            for (int k = 0; k < 3; ++k) {
                const float lo = blo3[k], hi = bhi3[k];
                h_lo[(size_t)o * 3 + k] = lo;
                // one scale for the three axes (cubic Morton cells): a tall block must not get tall cells
                h_inv[(size_t)o * 3 + k] = 1024.0f / (getenv("PTB_BVH_ANISO") ? std::max(hi - lo, 1e-30f) : ext);
                const float c = (0.5f * (lo + hi) - slo[k]) / std::max(shi[k] - slo[k], 1e-30f) * 1024.0f;
                code |= ex10((unsigned)std::min(std::max(c, 0.0f), 1023.0f)) << (2 - k);
            }
            order.push_back(std::make_pair(code, o));
        }
        std::sort(order.begin(), order.end());
        if (!flat) for (int r = 0; r < n_objs; ++r) h_rank[order[r].second] = r;
        cudaMemcpyAsync(obj_lo, h_lo.data(), 12 * (size_t)n_objs, cudaMemcpyHostToDevice, st);
        cudaMemcpyAsync(obj_inv, h_inv.data(), 12 * (size_t)n_objs, cudaMemcpyHostToDevice, st);
        cudaMemcpyAsync(rank, h_rank.data(), 4 * (size_t)n_objs, cudaMemcpyHostToDevice, st);
        k_morton<<<g, 256, 0, st>>>(centroid, d.tri_obj, n, obj_lo, obj_inv, keys[0], morton_orig, vals[0]);
        ++g_launches;
        // stable LSD radix sort (ptb_sort.cuh): 4 passes x 8 bits on the Morton code, then the object rank on top
        int cur = 0;
        int rank_passes = 0;
        if (!flat) for (unsigned m = (unsigned)(n_objs - 1); m; m >>= 8) ++rank_passes;
        for (int pass = 0; pass < 4 + rank_passes; ++pass) {
            if (pass == 4) { k_rank_keys<<<g, 256, 0, st>>>(vals[cur], d.tri_obj, rank, n, keys[cur]); ++g_launches; }
            const int shift = pass < 4 ? 8 * pass : 8 * (pass - 4);
            ps_sort_pass(keys[cur], (const unsigned *)vals[cur], d_n, shift, n_tiles, ghist, totals, keys[cur ^ 1], (unsigned *)vals[cur ^ 1], st);
            g_launches += 3;
            cur ^= 1;
        }
        if (cur == 0) cudaMemcpyAsync(vals[1], vals[0], 4 * (size_t)n, cudaMemcpyDeviceToDevice, st);
        k_key64<<<g, 256, 0, st>>>(vals[1], d.tri_obj, rank, morton_orig, n, key64);
        ++g_launches;
        cudaMemsetAsync(flags, 0, 4 * (size_t)n, st);
        k_karras<<<g, 256, 0, st>>>(key64, n, child, parent_inner, parent_leaf);
        k_refit<<<g, 256, 0, st>>>(n, vals[1], tri_box, child, parent_inner, parent_leaf, node_box, node_minref, flags, d.tri_pid, node_pid);
        k_pack<<<g, 256, 0, st>>>(n, vals[1], tri_box, child, node_box, node_minref, s->pad, nodes, d.tri_pid, node_pid);
        k_gather_tris<<<g, 256, 0, st>>>(n, vals[1], src_tri, tri_sorted);
        g_launches += 4;
    }
    cudaEventRecord(e1, st);
    cudaError_t err = cudaStreamSynchronize(st);
    if (err == cudaSuccess) err = cudaGetLastError();
    float ms = 0.f;
    if (err == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
    if (err != cudaSuccess) { set_error("LBVH build failed: %s", cudaGetErrorString(err)); return PTB_ERR_CUDA; }
    *nodes_out = nodes;
    *tri_out = tri_sorted;    // leaf order
    *ids_out = vals[1];
    *ms_out = ms;
    return PTB_OK;
}

extern "C" const char *ptb_last_error(void) { return g_err; }
namespace ptb { static int planes_optin_smem(); }     // ptb_api_render.inl
extern "C" int ptb_version(void) { return 100; }

// ---- replicated hierarchies (ptb_bvh_export / ptb_bvh_import): one device buffer = header + six sections
struct BvhBlobHeader {
    unsigned magic, version;
    int n_tri, n_tri_obj, has_occ, has_pid;
    float pad;
    unsigned long long off[7];      // nodes, tri, ids, nodes_occ, tri_occ, ids_occ, plane ids (bytes from the start of the buffer)
    unsigned long long total;
};
static const unsigned BVH_BLOB_MAGIC = 0x56425450u /* "PTBV" */, BVH_BLOB_VERSION = 2u;
static BvhBlobHeader bvh_blob_layout(int n_tri, int n_tri_obj, bool has_occ, bool has_pid, float pad) {
    BvhBlobHeader h;
    memset(&h, 0, sizeof(h));
    h.magic = BVH_BLOB_MAGIC; h.version = BVH_BLOB_VERSION;
    h.n_tri = n_tri; h.n_tri_obj = n_tri_obj; h.has_occ = has_occ ? 1 : 0; h.has_pid = has_pid ? 1 : 0; h.pad = pad;
    const size_t no = has_occ ? (size_t)n_tri_obj : 0;
    const size_t bytes[7] = {sizeof(BvhNode) * (size_t)n_tri, sizeof(TriRec) * (size_t)n_tri, 4 * (size_t)n_tri,
                             sizeof(BvhNode) * no, sizeof(TriRec) * no, 4 * no, has_pid ? (size_t)n_tri : 0};
    size_t at = 256;
    for (int i = 0; i < 7; ++i) { h.off[i] = at; at += (bytes[i] + 255) & ~(size_t)255; }
    h.total = at;
    return h;
}

static int scene_create_impl(const ptb_scene_desc *desc, const void *d_blob, size_t blob_bytes, void *stream, ptb_scene **out);

extern "C" int ptb_scene_create(const ptb_scene_desc *desc, void *stream, ptb_scene **out) {
    return scene_create_impl(desc, nullptr, 0, stream, out);
}

extern "C" int ptb_bvh_import(const ptb_scene_desc *desc, const void *d_blob, size_t blob_bytes, void *stream, ptb_scene **out) {
    if (!d_blob || blob_bytes < 256) { set_error("ptb_bvh_import: no hierarchy buffer"); return PTB_ERR_ARG; }
    return scene_create_impl(desc, d_blob, blob_bytes, stream, out);
}

struct PhaseTimer {     // PTB_TIMING=1: host wall-clock of the phases of scene creation on stderr
    bool on; std::chrono::steady_clock::time_point t0;
    PhaseTimer() : on(getenv("PTB_TIMING") != nullptr), t0(std::chrono::steady_clock::now()) {}
    void mark(const char *what) {
        if (!on) return;
        const auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[ptb timing] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

static int scene_create_impl(const ptb_scene_desc *desc, const void *d_blob, size_t blob_bytes, void *stream, ptb_scene **out) {
    PhaseTimer pt_;
    if (!desc || !out) { set_error("ptb_scene_create: null argument"); return PTB_ERR_ARG; }
    if (desc->n_tri <= 0 || desc->n_tri_obj < 0 || desc->n_tri_obj > desc->n_tri || desc->n_obj <= 0 ||
        desc->n_light_tri != desc->n_tri - desc->n_tri_obj || desc->n_light_tri <= 0 || !desc->verts ||
        !desc->normals || !desc->tri_obj || !desc->mat || !desc->light_areas) {
        set_error("ptb_scene_create: inconsistent scene description");
        return PTB_ERR_ARG;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int accel = d_blob ? (int)PTB_ACCEL_LBVH : desc->accel;
    // measured on random soups (scratch/soup_speed.py): the plane list in shared memory wins up to ~256 triangles
    // (2.0 vs 1.4 Gsamples/s at 64, 255 vs 207 Msamples/s at 256), the LBVH from ~400 on (119 vs 163 at 512)
    if (accel == PTB_ACCEL_AUTO) accel = desc->n_tri <= PTB_BRUTE_AUTO_TRI ? PTB_ACCEL_BRUTE : PTB_ACCEL_LBVH;
    if (accel == PTB_ACCEL_BRUTE && desc->n_tri > PTB_BRUTE_MAX_TRI) {
        set_error("brute-force traversal holds at most %d triangles in shared memory", PTB_BRUTE_MAX_TRI);
        return PTB_ERR_UNSUPPORTED;
    }
    if (accel == PTB_ACCEL_LBVH && desc->n_tri < 2) { set_error("LBVH needs at least 2 triangles"); return PTB_ERR_UNSUPPORTED; }

    ptb_scene *s = new ptb_scene();
    memset(&s->d, 0, sizeof(DScene));
    cudaGetDevice(&s->device);
    s->create_stream = (cudaStream_t)stream;
    s->streams.push_back(s->create_stream);
    const int n = desc->n_tri, no = desc->n_obj;
    // Large scenes (LBVH): the FP64 arrays are uploaded straight from the caller's buffers and the FP32 records,
    // shading normals, scene extent and scale are computed on the device (ptb_bvh_build.cuh); small scenes keep
    // the host loop because the plane grouping below needs the records on the host.
    const bool dev_records = accel == PTB_ACCEL_LBVH;
    const bool dev_arrays = (desc->reserved & PTB_SCENE_ARRAYS_ON_DEVICE) != 0;      // verts / normals / tri_obj are device pointers
    if (dev_arrays && !dev_records) { delete s; set_error("PTB_SCENE_ARRAYS_ON_DEVICE needs an LBVH scene (small scenes are grouped into planes on the host)"); return PTB_ERR_UNSUPPORTED; }
    const cudaMemcpyKind up = dev_arrays ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    DScene &d = s->d;
    int rc = PTB_OK;
    double scale = 1.0, blo[3] = {1e300, 1e300, 1e300}, bhi[3] = {-1e300, -1e300, -1e300};
    std::vector<TriRec> tri;
    std::vector<float4> sn;
    std::vector<int> ids;
    int ext_blocks = 0;
    double *d_ext = nullptr;
    // page-locked landing buffer for the extent pass: a device-to-host copy into pageable memory would block the host
    // until the uploads ahead of it in the stream are done, and the host has tables to compute meanwhile
    struct Pinned { double *p = nullptr; size_t cap = 0; ~Pinned() { if (p) cudaFreeHost(p); } };
    static thread_local Pinned pinned;       // one per host thread: scenes may be created concurrently
    double *&h_ext_pinned = pinned.p; size_t &h_ext_cap = pinned.cap;
    struct ExtView { double *p; size_t n; double *data() const { return p; } size_t size() const { return n; } void resize(size_t) {} } h_ext = {nullptr, 0};
    if (dev_records) {      // issue the big uploads and the extent pass; the host tables below are computed while they run
        if ((rc = upload_raw(s, desc->verts, (size_t)n * 9, &d.verts64, st, up)) || (rc = upload_raw(s, desc->normals, (size_t)n * 3, &d.normals64, st, up)) ||
            (rc = upload_raw(s, desc->tri_obj, (size_t)n, &d.tri_obj, st, up))) { ptb_scene_destroy(s); return rc; }
        ext_blocks = grid_for((size_t)n * 3, 256, 4);
        h_ext.n = (size_t)ext_blocks * 8;
        if (h_ext_cap < h_ext.n) {
            if (h_ext_pinned) cudaFreeHost(h_ext_pinned);
            h_ext_pinned = nullptr; h_ext_cap = 0;
            if (cudaMallocHost((void **)&h_ext_pinned, h_ext.n * sizeof(double)) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of page-locked host memory"); return PTB_ERR_CUDA; }
            h_ext_cap = h_ext.n;
        }
        h_ext.p = h_ext_pinned;
        if (scene_malloc((void **)&d_ext, h_ext.size() * sizeof(double), st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory"); return PTB_ERR_CUDA; }
        k_scene_extent<<<ext_blocks, 256, 0, st>>>(d.verts64, (size_t)n * 3, d.tri_obj, n, no, d_ext);
        ++g_launches;
        cudaMemcpyAsync(h_ext.data(), d_ext, h_ext.size() * sizeof(double), cudaMemcpyDeviceToHost, st);
    }
    pt_.mark("uploads issued");
    const int32_t *tobj = desc->tri_obj;
    std::vector<double> mat64((size_t)(no + 1) * 8, 0.0);
    std::copy(desc->mat, desc->mat + (size_t)no * 8, mat64.begin());
    std::vector<float> mat32(mat64.begin(), mat64.end());
    std::vector<float4> lv(dev_arrays ? 0 : (size_t)desc->n_light_tri * 3);
    for (int i = 0; i < desc->n_light_tri * 3 && !dev_arrays; ++i) {
        const double *p = desc->verts + 9 * (size_t)desc->n_tri_obj + 3 * (size_t)i;
        lv[i] = make_float4((float)p[0], (float)p[1], (float)p[2], 0.f);
    }
    std::vector<double> cdf(desc->n_light_tri + 1);
    double acc = 0.0, total = 0.0;
    cdf[0] = 0.0;
    for (int i = 0; i < desc->n_light_tri; ++i) {       // utils.py:30-35: sequential FP64 sums
        total = total + desc->light_areas[i];
        acc = acc + desc->light_areas[i];
        cdf[i + 1] = acc;
    }
    // Integer thresholds on the 24-bit uniform m = x >> 8 (u = m * 2^-24) that reproduce the reference's FP64
    // comparisons exactly: both products below are monotone in m, so a bisection finds the switch point.
    auto first_m = [](auto pred) {            // smallest m in [0, 2^24] with pred(m) true (pred monotone), 2^24 if none
        unsigned lo = 0, hi = 1u << 24;
        if (pred(0u)) return 0u;
        while (hi - lo > 1) { unsigned mid = (lo + hi) >> 1; if (pred(mid)) hi = mid; else lo = mid; }
        return pred(hi) ? hi : (1u << 24);
    };
    std::vector<unsigned> light_thr(desc->n_light_tri);
    for (int i = 0; i < desc->n_light_tri; ++i) {
        const double c = cdf[i];              // pick_random_triangle: cum[i] <= uniform(0, total)  (utils.py:30, 38)
        // smallest m with total * (m * 2^-24) >= c: start from the real-number estimate and walk to the switch point
        // (the product is monotone in m), a couple of evaluations instead of a 24-step bisection per light triangle
        auto reached = [&](long long m) { return 0.0 + (total - 0.0) * ((double)m * 5.9604644775390625e-08) >= c; };
        long long m;
        const double guess = c / total * 16777216.0;
        if (!(total > 0.0) || !(guess >= 0.0) || !(guess <= 16777216.0)) m = reached(0) ? 0 : (reached(1ll << 24) ? -1 : 1ll << 24);
        else m = (long long)guess;
        if (m < 0) m = first_m([&](unsigned mm) { return reached(mm); });          // degenerate areas: plain bisection
        else {
            while (m > 0 && reached(m - 1)) --m;
            while (m < (1ll << 24) && !reached(m)) ++m;
        }
        light_thr[i] = i == 0 ? 0u : (unsigned)m;
    }
    // bucket table over m: with B >= n_light_tri buckets a pick needs 1-2 threshold reads instead of log2(n)
    int light_shift = 24;
    while (light_shift > 2 && (1 << (24 - light_shift)) < desc->n_light_tri) --light_shift;
    const int n_bucket = 1 << (24 - light_shift);
    std::vector<int> light_bucket((size_t)n_bucket + 1);
    for (int b = 0, i = 0; b <= n_bucket; ++b) {            // largest i with thr[i] <= b << shift (thr is non-decreasing, thr[0] = 0)
        const unsigned long long edge = (unsigned long long)b << light_shift;
        while (i + 1 < desc->n_light_tri && light_thr[i + 1] <= edge) ++i;
        light_bucket[b] = i;
    }
    d.light_shift = light_shift;
    std::vector<int> lobe_thr(no + 1, 0), pow_int(no + 1, -1);
    for (int k = 0; k < no; ++k) {
        const double kd = desc->mat[8 * k + 4], ks = desc->mat[8 * k + 5], ne = desc->mat[8 * k + 7];
        // diffuse iff uniform(0, kd + ks) <= kd  (main.py:240-241)  <=>  m < first m with the product > kd
        lobe_thr[k] = (int)first_m([&](unsigned m) { return !(0.0 + ((kd + ks) - 0.0) * ((double)m * 5.9604644775390625e-08) <= kd); });
        if (ne >= 0.0 && ne <= 64.0 && ne == floor(ne)) pow_int[k] = (int)ne;
    }
    std::vector<unsigned long long> ctr(CTR_COUNT, 0ull);
    std::vector<int> obj_first(no + 2, n);
    if (!dev_arrays) {
        for (int i = n - 1; i >= 0; --i) { const int o = tobj[i]; if ((unsigned)o <= (unsigned)no) obj_first[o] = i; }   // range is checked below
        for (int k = no; k >= 0; --k) if (obj_first[k] == n && k + 1 <= no + 1) obj_first[k] = obj_first[k + 1];
    }
    pt_.mark("light tables, obj_first (host)");
    if (dev_records) {
        cudaError_t e = cudaStreamSynchronize(st);
        cudaFreeAsync(d_ext, st);
        if (e != cudaSuccess) { ptb_scene_destroy(s); set_error("scene extent: %s", cudaGetErrorString(e)); return PTB_ERR_CUDA; }
        bool bad_obj = false;
        for (int b = 0; b < ext_blocks; ++b) {
            const double *x = h_ext.data() + 8 * (size_t)b;
            for (int k = 0; k < 3; ++k) { blo[k] = std::min(blo[k], x[k]); bhi[k] = std::max(bhi[k], x[3 + k]); }
            scale = std::max(scale, x[6]);
            bad_obj = bad_obj || x[7] != 0.0;
        }
        if (bad_obj) { ptb_scene_destroy(s); set_error("tri_obj out of range"); return PTB_ERR_ARG; }
    } else {
        for (size_t i = 0; i < (size_t)n * 9; ++i) scale = std::max(scale, fabs(desc->verts[i]));
        for (size_t i = 0; i < (size_t)n * 3; ++i)
            for (int k = 0; k < 3; ++k) { blo[k] = std::min(blo[k], desc->verts[3 * i + k]); bhi[k] = std::max(bhi[k], desc->verts[3 * i + k]); }
    }
    pt_.mark("upload + extent (wait)");
    for (int i = 0; i < 3; ++i) scale = std::max(scale, fabs(desc->eye[i]));
    s->dt_scale = (float)(32.0 * 5.9604644775390625e-08 * scale);
    for (int k = 0; k < 3; ++k) { d.bb_lo[k] = (float)blo[k]; d.bb_inv[k] = (float)(256.0 / std::max(bhi[k] - blo[k], 1e-30)); }

    if (dev_records) {
        TriRec *d_tri = nullptr; float4 *d_sn = nullptr;
        if (scene_malloc((void **)&d_tri, sizeof(TriRec) * (size_t)n, st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory"); return PTB_ERR_CUDA; }
        s->allocs.push_back(d_tri);
        if (scene_malloc((void **)&d_sn, sizeof(float4) * (size_t)n, st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory"); return PTB_ERR_CUDA; }
        s->allocs.push_back(d_sn);
        k_make_records<<<grid_for((size_t)n, 256, 8), 256, 0, st>>>(d.verts64, d.normals64, d.tri_obj, n, scale, d_tri, d_sn);
        ++g_launches;
        d.tri = d_tri; d.tri_sn = d_sn; d.tri_id = nullptr;
        // supporting-plane ids (ptb_bvh_build.cuh): a received hierarchy brings its own numbering (ptb_bvh_import)
        if (!d_blob && getenv("PTB_NO_PLANE_CULL") == nullptr) {
            unsigned char *d_pid = nullptr; unsigned long long *d_table = nullptr;
            if (scene_malloc((void **)&d_pid, (size_t)n, st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory"); return PTB_ERR_CUDA; }
            s->allocs.push_back(d_pid);
            if (scene_malloc((void **)&d_table, 8 * (PTB_PLANE_IDS + 2), st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory"); return PTB_ERR_CUDA; }
            s->allocs.push_back(d_table);          // 2 KB, lives with the handle: no error path can leak it
            cudaMemsetAsync(d_table, 0, 8 * (PTB_PLANE_IDS + 2), st);
            // normals within q_n and offsets within q_d: q_n * diameter and q_d stay 4 x below sqrt(1e-5) * 1e-5 = 3.2e-8
            double diam = 0.0;
            for (int k = 0; k < 3; ++k) diam = std::max(diam, bhi[k] - blo[k]);
            const double q_n = std::min(2e-10, 8e-9 / (2.0 * std::max(diam, 1e-30))), q_d = 8e-9;
            k_plane_ids<<<grid_for((size_t)n, 256, 8), 256, 0, st>>>(d.verts64, n, q_n, q_d, d_table, d_pid);
            ++g_launches;
            std::vector<unsigned long long> h_table(PTB_PLANE_IDS + 2);
            cudaMemcpyAsync(h_table.data(), d_table, 8 * (PTB_PLANE_IDS + 2), cudaMemcpyDeviceToHost, st);
            cudaError_t e = cudaStreamSynchronize(st);
            if (e != cudaSuccess) { ptb_scene_destroy(s); set_error("plane ids: %s", cudaGetErrorString(e)); return PTB_ERR_CUDA; }
            const bool overflow = h_table[PTB_PLANE_IDS + 1] != 0;
            if (!overflow) {
                // the table's slots were handed out in arrival order; renumber the planes by the rank of their keys so
                // that the ids (they travel in the nodes and in exported hierarchies) do not depend on the race
                std::vector<unsigned long long> keys_sorted;
                for (int i = 0; i < PTB_PLANE_IDS; ++i) if (h_table[i]) keys_sorted.push_back(h_table[i]);
                std::sort(keys_sorted.begin(), keys_sorted.end());
                unsigned char remap[256];
                memset(remap, 0, sizeof(remap));
                for (int i = 0; i < PTB_PLANE_IDS; ++i)
                    if (h_table[i]) remap[i + 1] = (unsigned char)(1 + (std::lower_bound(keys_sorted.begin(), keys_sorted.end(), h_table[i]) - keys_sorted.begin()));
                // the table's device memory is free to carry the 256-byte map back (stream-ordered after the read)
                cudaMemcpyAsync(d_table, remap, 256, cudaMemcpyHostToDevice, st);
                k_remap_pid<<<grid_for((size_t)n, 256, 8), 256, 0, st>>>(d_pid, n, (const unsigned char *)d_table);
                ++g_launches;
                cudaStreamSynchronize(st);        // `remap` is on the stack
            }
            d.tri_pid = overflow ? nullptr : d_pid;      // more planes than ids (a soup): no plane culling for this scene
        }
    } else {
        tri.resize(n); sn.resize(n); ids.resize(n);
        for (int i = 0; i < n; ++i) {
            if (desc->tri_obj[i] < 0 || desc->tri_obj[i] > no) { delete s; set_error("tri_obj[%d] out of range", i); return PTB_ERR_ARG; }
            tri[i] = make_trirec(desc->verts + 9 * (size_t)i, desc->tri_obj[i], scale);
            tri[i].q3.w = bits_f(i);                     // reference index travels with the record (LBVH leaf order)
            const double *nn = desc->normals + 3 * (size_t)i;
            sn[i] = make_float4((float)nn[0], (float)nn[1], (float)nn[2], bits_f(desc->tri_obj[i]));
            ids[i] = i;
        }
    }
    pt_.mark("records");
    std::vector<PlaneRec> planes; int n_occ = 0;
    if (accel == PTB_ACCEL_BRUTE) build_plane_groups(desc, tri, planes, n_occ);

    if (!dev_records &&
        ((rc = upload(s, tri, &d.tri, st)) || (rc = upload(s, ids, &d.tri_id, st)) || (rc = upload(s, sn, &d.tri_sn, st)) ||
         (rc = upload_raw(s, desc->verts, (size_t)n * 9, &d.verts64, st)) || (rc = upload_raw(s, desc->normals, (size_t)n * 3, &d.normals64, st)) ||
         (rc = upload_raw(s, desc->tri_obj, (size_t)n, &d.tri_obj, st)))) {
        ptb_scene_destroy(s);
        return rc;
    }
    if ((rc = upload(s, mat32, &d.mat32, st)) || (rc = upload(s, lv, &d.light_v32, st)) ||
        (rc = upload(s, mat64, &d.mat64, st)) || (rc = upload(s, cdf, &d.light_cdf, st)) ||
        (rc = upload(s, obj_first, &d.obj_first, st)) ||
        (rc = upload(s, light_thr, &d.light_thr, st)) || (rc = upload(s, light_bucket, &d.light_bucket, st)) || (rc = upload(s, lobe_thr, &d.obj_lobe_thr, st)) || (rc = upload(s, pow_int, &d.obj_pow_int, st)) || (rc = upload(s, planes, &s->planes.planes, st)) || (rc = upload(s, ctr, (const unsigned long long **)&d.counters, st))) {
        ptb_scene_destroy(s);
        return rc;
    }
    if (dev_arrays) {       // the two tables the host could not derive: computed from the device arrays (ptb_bvh_build.cuh)
        float4 *d_lv = nullptr;
        if (scene_malloc((void **)&d_lv, sizeof(float4) * 3 * (size_t)desc->n_light_tri, st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory"); return PTB_ERR_CUDA; }
        s->allocs.push_back(d_lv);
        k_light_verts32<<<grid_for(3 * (size_t)desc->n_light_tri, 256, 8), 256, 0, st>>>(d.verts64, desc->n_tri_obj, desc->n_light_tri, d_lv);
        d.light_v32 = d_lv;
        int *d_first = const_cast<int *>(d.obj_first);          // uploaded above, every entry = n
        k_obj_first<<<grid_for((size_t)n, 256, 8), 256, 0, st>>>(d.tri_obj, n, no, d_first);
        k_obj_first_fill<<<1, 32, 0, st>>>(d_first, no, n);
        g_launches += 3;
    }
    pt_.mark("small uploads");
    d.n_tri = n; d.n_tri_obj = desc->n_tri_obj; d.n_obj = no; d.n_light_tri = desc->n_light_tri;
    d.accel = accel;
    for (int i = 0; i < 3; ++i) { d.eye[i] = desc->eye[i]; d.light_color[i] = desc->light_color[i]; }
    for (int i = 0; i < 4; ++i) d.ortho[i] = desc->ortho[i];
    d.ambient = desc->ambient;
    d.light_total = total;
    s->planes.n_planes = (int)planes.size(); s->planes.n_occ = n_occ;
    if (accel == PTB_ACCEL_BRUTE && (rc = planes_optin_smem())) { ptb_scene_destroy(s); return rc; }
    s->use_planes = getenv("PTB_BRUTE_V1") == nullptr;
    { const char *x2 = getenv("PTB_FUSED_X2"); s->fused_x2 = !(x2 && x2[0] == '0'); }
    s->bvh_dynamic = getenv("PTB_BVH_STATIC") == nullptr;
    s->sort_rays = getenv("PTB_NO_SORT") == nullptr;
    s->build_ms = 0.0;
    s->n_nodes = 0;
    s->pad = (float)(scale * (getenv("PTB_BVH_PAD18") ? 3.814697265625e-06 : 4.76837158203125e-07));   // 2^-21 * scene scale (4 x the FP32 slab error)
    if (accel == PTB_ACCEL_LBVH) {
        const TriRec *ref_order = d.tri;      // the reference-order array stays allocated: source of both gathers
        double ms_all = 0.0, ms_occ = 0.0;
        if (d_blob) {
            // received hierarchy: check the header against this scene, then six device-to-device copies
            BvhBlobHeader h;
            if (cudaMemcpyAsync(&h, d_blob, sizeof(h), cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
                ptb_scene_destroy(s); set_error("ptb_bvh_import: cannot read the buffer header"); return PTB_ERR_CUDA;
            }
            const bool want_occ = d.n_tri_obj >= 2;
            const BvhBlobHeader w = bvh_blob_layout(n, d.n_tri_obj, want_occ, h.has_pid == 1, s->pad);
            if (h.magic != BVH_BLOB_MAGIC || h.version != BVH_BLOB_VERSION || h.n_tri != n || h.n_tri_obj != d.n_tri_obj ||
                h.has_occ != w.has_occ || (h.has_pid != 0 && h.has_pid != 1) || h.pad != s->pad || h.total != w.total || blob_bytes < w.total) {
                ptb_scene_destroy(s);
                set_error("ptb_bvh_import: buffer does not belong to this scene (magic %08x version %u, %d/%d triangles, %llu bytes; expected %d/%d, %llu)",
                          h.magic, h.version, h.n_tri, h.n_tri_obj, (unsigned long long)h.total, n, d.n_tri_obj, (unsigned long long)w.total);
                return PTB_ERR_ARG;
            }
            void *sec[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
            for (int i = 0; i < 7; ++i) {
                const size_t bytes = (size_t)((i + 1 < 7 ? w.off[i + 1] : w.total) - w.off[i]);
                if (bytes == 0) continue;          // no occlusion tree / no plane ids
                if (scene_malloc(&sec[i], bytes, st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory"); return PTB_ERR_CUDA; }
                s->allocs.push_back(sec[i]);
                if (cudaMemcpyAsync(sec[i], (const char *)d_blob + w.off[i], bytes, cudaMemcpyDeviceToDevice, st) != cudaSuccess) {
                    ptb_scene_destroy(s); set_error("ptb_bvh_import: copy failed"); return PTB_ERR_CUDA;
                }
            }
            d.nodes = (const BvhNode *)sec[0]; d.tri = (const TriRec *)sec[1]; d.tri_id = (const int *)sec[2];
            d.nodes_occ = (const BvhNode *)sec[3]; d.tri_occ = (const TriRec *)sec[4]; d.tri_id_occ = (const int *)sec[5];
            d.tri_pid = (const unsigned char *)sec[6];
            s->build_ms = 0.0;
            s->n_nodes = n - 1;
        } else {
        void *arena = nullptr;
        const size_t arena_bytes = lbvh_arena_bytes(n, no + 1);
        if (scene_malloc(&arena, arena_bytes, st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory (LBVH build arena, %zu bytes)", arena_bytes); return PTB_ERR_CUDA; }
        if ((rc = build_lbvh(s, n, no + 1, ref_order, &d.nodes, &d.tri, &d.tri_id, &ms_all, arena, arena_bytes, st))) { cudaFreeAsync(arena, st); ptb_scene_destroy(s); return rc; }
        if (d.n_tri_obj >= 2) {
            if ((rc = build_lbvh(s, d.n_tri_obj, no, ref_order, &d.nodes_occ, &d.tri_occ, &d.tri_id_occ, &ms_occ, arena, arena_bytes, st))) { cudaFreeAsync(arena, st); ptb_scene_destroy(s); return rc; }
            cudaFreeAsync(arena, st);
        } else {
            cudaFreeAsync(arena, st);                              // degenerate occlusion set: fall back to the full tree + min-ref pruning
            d.nodes_occ = nullptr; d.tri_occ = nullptr; d.tri_id_occ = nullptr;
        }
        s->build_ms = ms_all + ms_occ;
        s->n_nodes = n - 1;
        }
    }
#if PTB_TOP_K > 0
    if (accel == PTB_ACCEL_LBVH && n < (1 << 24)) {      // reference indices share a word with the plane ids in a TopNode
        // breadth-first tops of both trees, staged in shared memory by every block of the traversal kernels
        TopNode *tops = nullptr; int *d_cnt = nullptr;
        if (scene_malloc((void **)&tops, 2 * (size_t)PTB_TOP_K * sizeof(TopNode) + 64, st) != cudaSuccess) { ptb_scene_destroy(s); set_error("out of device memory"); return PTB_ERR_CUDA; }
        s->allocs.push_back(tops);
        d_cnt = (int *)(tops + 2 * PTB_TOP_K);
        int h_cnt[2] = {0, 0};
        k_build_top<<<1, 32, 0, st>>>(d.nodes, PTB_TOP_K, tops, d_cnt);
        if (d.nodes_occ) k_build_top<<<1, 32, 0, st>>>(d.nodes_occ, PTB_TOP_K, tops + PTB_TOP_K, d_cnt + 1);
        g_launches += d.nodes_occ ? 2 : 1;
        cudaMemcpyAsync(h_cnt, d_cnt, sizeof(h_cnt), cudaMemcpyDeviceToHost, st);
        if (cudaStreamSynchronize(st) != cudaSuccess) { ptb_scene_destroy(s); set_error("top-of-tree build failed: %s", cudaGetErrorString(cudaGetLastError())); return PTB_ERR_CUDA; }
        d.top = tops; d.top_n = h_cnt[0];
        d.top_occ = d.nodes_occ ? tops + PTB_TOP_K : nullptr; d.top_occ_n = h_cnt[1];
    }
#endif
    pt_.mark("hierarchies");
    PTB_CUDA(cudaStreamSynchronize(st));   // host staging vectors go out of scope
    *out = s;
    return PTB_OK;
}

extern "C" int ptb_scene_destroy(ptb_scene *s) {
    if (!s) return PTB_OK;
    // every stream that used the scene must be done with it before its memory goes back to the pool: wait for the
    // streams the handle has seen (not for the whole device - other work of the process keeps running), then free in
    // the order of the stream the scene was created on
    int cur = 0;
    cudaGetDevice(&cur);
    if (cur != s->device) cudaSetDevice(s->device);
    for (cudaStream_t st : s->streams) cudaStreamSynchronize(st);
    for (void *p : s->allocs) cudaFreeAsync(p, s->create_stream);
    if (cur != s->device) cudaSetDevice(cur);
    delete s;
    return PTB_OK;
}

extern "C" int ptb_scene_info(const ptb_scene *s, int32_t *accel, int32_t *n_nodes, double *build_ms) {
    if (!s) { set_error("null scene"); return PTB_ERR_ARG; }
    if (accel) *accel = s->d.accel;
    if (n_nodes) *n_nodes = s->n_nodes;
    if (build_ms) *build_ms = s->build_ms;
    return PTB_OK;
}

extern "C" size_t ptb_bvh_export_bytes(const ptb_scene *s) {
    if (!s || s->d.accel != PTB_ACCEL_LBVH || !s->d.nodes) return 0;
    return (size_t)bvh_blob_layout(s->d.n_tri, s->d.n_tri_obj, s->d.nodes_occ != nullptr, s->d.tri_pid != nullptr, s->pad).total;
}

extern "C" int ptb_bvh_export(const ptb_scene *s, void *d_blob, size_t blob_bytes, void *stream) {
    if (!s || !d_blob) { set_error("ptb_bvh_export: null argument"); return PTB_ERR_ARG; }
    if (s->d.accel != PTB_ACCEL_LBVH || !s->d.nodes) { set_error("ptb_bvh_export: the scene has no LBVH (shared-memory plane list)"); return PTB_ERR_UNSUPPORTED; }
    const DScene &d = s->d;
    const BvhBlobHeader h = bvh_blob_layout(d.n_tri, d.n_tri_obj, d.nodes_occ != nullptr, d.tri_pid != nullptr, s->pad);
    if (blob_bytes < h.total) { set_error("ptb_bvh_export: buffer of %zu bytes, %llu needed", blob_bytes, (unsigned long long)h.total); return PTB_ERR_NOMEM; }
    cudaStream_t st = (cudaStream_t)stream;
    // zero the header block and the alignment gaps so equal scenes export equal buffers
    PTB_CUDA(cudaMemsetAsync(d_blob, 0, 256, st));
    PTB_CUDA(cudaMemcpyAsync(d_blob, &h, sizeof(h), cudaMemcpyHostToDevice, st));
    const void *src[7] = {d.nodes, d.tri, d.tri_id, d.nodes_occ, d.tri_occ, d.tri_id_occ, d.tri_pid};
    const size_t no = h.has_occ ? (size_t)d.n_tri_obj : 0;
    // a tree over n leaves has n - 1 nodes; the n-th slot of the node array is never written by the build
    const size_t used[7] = {sizeof(BvhNode) * (size_t)(d.n_tri - 1), sizeof(TriRec) * (size_t)d.n_tri, 4 * (size_t)d.n_tri,
                            no ? sizeof(BvhNode) * (no - 1) : 0, sizeof(TriRec) * no, 4 * no, d.tri_pid ? (size_t)d.n_tri : 0};
    for (int i = 0; i < 7; ++i) {
        const size_t span = (size_t)((i + 1 < 7 ? h.off[i + 1] : h.total) - h.off[i]);
        if (used[i]) PTB_CUDA(cudaMemcpyAsync((char *)d_blob + h.off[i], src[i], used[i], cudaMemcpyDeviceToDevice, st));
        if (span > used[i]) PTB_CUDA(cudaMemsetAsync((char *)d_blob + h.off[i] + used[i], 0, span - used[i], st));
    }
    PTB_CUDA(cudaStreamSynchronize(st));      // the header came from the stack
    return PTB_OK;
}

extern "C" int64_t ptb_pixel_count(const ptb_pixels *px) { return px ? pixel_count(px) : 0; }

#include "ptb_api_render.inl"
#include "ptb_obj.inl"

// ---------------------------------------------------------------- ptb_trace_closest without any allocation
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  ptb_trace_closest takes its two ray queues and the count word from cudaMallocAsync; this variant takes them
// from a file-scope device array of a fixed, stated size.  Every GPU that loads the library holds the array until
// exit.  In HOST code the array's name is only a stand-in: the device address is asked for once per device with
// cudaGetSymbolAddress (that device current) and it is that pointer the kernels get.
#define PTB_NOALLOC_SCRATCH_BYTES (2 * sizeof(double4) * (size_t)PTB_TRACE_CLOSEST_NOALLOC_MAX_RAYS + 256)
static __device__ __align__(256) unsigned char g_noalloc_scratch[PTB_NOALLOC_SCRATCH_BYTES];

namespace ptb {
static void *g_noalloc_ptr[PTB_MAX_DEVICES];        // device address of g_noalloc_scratch on device i, guarded by g_state_mu
static int noalloc_scratch(char **out) {
    int dev = 0;
    PTB_CUDA(cudaGetDevice(&dev));
    if ((unsigned)dev >= PTB_MAX_DEVICES) { set_error("device %d: more devices than the library keeps state for", dev); return PTB_ERR_UNSUPPORTED; }
    std::lock_guard<std::mutex> lk(g_state_mu);
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        PTB_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) { set_error("cudaGetSymbolAddress gave no address for the scratch array on device %d", dev); return PTB_ERR_CUDA; }
        g_noalloc_ptr[dev] = p;
    }
    *out = (char *)g_noalloc_ptr[dev];
    return PTB_OK;
}
}  // namespace ptb

extern "C" int ptb_trace_closest_noalloc(const ptb_scene *s, const void *d_rays, int64_t n, int32_t precision,
                                         int32_t *d_tri, void *d_t, void *stream) {
    int rc;
    if ((rc = check_scene(s, stream))) return rc;
    if (!d_rays || !d_tri || !d_t || n < 0) { set_error("ptb_trace_closest_noalloc: bad argument"); return PTB_ERR_ARG; }
    if (precision != PTB_PREC_FP32 && precision != PTB_PREC_FP64) { set_error("unknown precision %d", precision); return PTB_ERR_ARG; }
    if (n > PTB_TRACE_CLOSEST_NOALLOC_MAX_RAYS) {
        set_error("ptb_trace_closest_noalloc: %lld rays, the fixed scratch array holds at most %d per call (split the call, or use ptb_trace_closest)",
                  (long long)n, PTB_TRACE_CLOSEST_NOALLOC_MAX_RAYS);
        return PTB_ERR_NOMEM;
    }
    if (n == 0) return PTB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    char *scratch = nullptr;
    if ((rc = noalloc_scratch(&scratch))) return rc;
    // The array keeps the last call's contents.  Both queues are rewritten for [0, n) before they are read and nothing
    // beyond n is read; the count word is the one thing a kernel reads that must be right from the start: cleared here,
    // then set by k_rays_to_queue*.
    const size_t v = precision == PTB_PREC_FP64 ? sizeof(double4) : sizeof(float4);
    unsigned *cnt = (unsigned *)(scratch + 2 * v * (size_t)n);
    PTB_CUDA(cudaMemsetAsync(cnt, 0, 256, st));
    if (precision == PTB_PREC_FP32) {
        float4 *qo = (float4 *)scratch, *qd = qo + n;
        { ProfScope ps_(PK_OTHER, st);
        k_rays_to_queue32<<<grid_for((size_t)n, 256, 8), 256, 0, st>>>((const float *)d_rays, n, qo, qd, cnt);
        }
        PTB_LAUNCHED();
        Queues<float> q; memset(&q, 0, sizeof(q));
        q.qo[0] = qo; q.qd[0] = qd;
        if ((rc = launch_closest32(s, q, 0, cnt, (size_t)n, false, (float *)d_t, d_tri, st))) return rc;
    } else {
        double4 *qo = (double4 *)scratch, *qd = qo + n;
        { ProfScope ps_(PK_OTHER, st);
        k_rays_to_queue64<<<grid_for((size_t)n, 256, 8), 256, 0, st>>>((const double *)d_rays, n, qo, qd, cnt);
        }
        PTB_LAUNCHED();
        if ((rc = launch_closest64(s, qo, qd, cnt, (size_t)n, (double *)d_t, d_tri, st))) return rc;
    }
    return PTB_OK;
}
