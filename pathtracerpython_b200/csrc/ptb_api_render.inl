// ptb_api_render.inl - included by ptb_api.cu: the launch sequences behind ptb_trace_primary,
// ptb_trace_closest, ptb_trace_paths, ptb_tonemap, ptb_scatter_accum and the counters.

namespace ptb {

// The plane list of a 512-triangle soup (up to 512 records of 96 B) exceeds the 48 KB a kernel gets without asking
static int planes_optin_smem() {
    DeviceState *ds = device_state();      // cudaFuncSetAttribute is per device
    if (ds->planes_optin.load(std::memory_order_acquire)) return PTB_OK;
    std::lock_guard<std::mutex> lk(g_state_mu);
    const int bytes = (int)(PTB_BRUTE_MAX_TRI * sizeof(PlaneRec) + 16);
    PTB_CUDA(cudaFuncSetAttribute(k_closest_planes32<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute(k_closest_planes32<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute(k_shadow_planes32, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute(k_shadow_planes_phys32, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute(k_bounce_planes32<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute(k_bounce_planes32<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute((k_bounce_planes32x2<true, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute((k_bounce_planes32x2<true, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute((k_bounce_planes32x2<false, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    PTB_CUDA(cudaFuncSetAttribute((k_bounce_planes32x2<false, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    ds->planes_optin.store(true, std::memory_order_release);
    return PTB_OK;
}

// FP64 re-trace of uncertain FP32 primary rays (bounce 0: queue position == path id).
__global__ void __launch_bounds__(128)
k_retrace64(DScene sc, PixelMap pm, int npix, const unsigned *__restrict__ unc_count, const int *__restrict__ unc_list,
            float *__restrict__ hit_t, int *__restrict__ hit_tri) {
    const unsigned count = *unc_count;
    for (unsigned u = blockIdx.x * blockDim.x + threadIdx.x; u < count; u += gridDim.x * blockDim.x) {
        const int i = unc_list[u];
        const int r = pm.ray_index(i % npix);
        d3 o, d;
        primary_ray64(sc, pm.W, pm.H, r, o, d);
        const Hit64 h = sc.accel == PTB_ACCEL_LBVH ? closest_bvh64(sc, o, d) : closest_brute64(sc, o, d);
        hit_tri[i] = h.tri;
        if (h.tri >= 0) {
            const d3 vr = xdiv3(d, xnorm(d));
            hit_t[i] = (float)xdot(xsub3(h.P, o), vr);
        } else {
            hit_t[i] = 0.f;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(sc.counters + CTR_RETRACED, (unsigned long long)count);
}

__global__ void k_rays_to_queue32(const float *__restrict__ rays, long long n, float4 *qo, float4 *qd, unsigned *count) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float *r = rays + 6 * i;
        const float inv = rsqrtf(r[3] * r[3] + r[4] * r[4] + r[5] * r[5]);
        qo[i] = make_float4(r[0], r[1], r[2], 1.f);
        qd[i] = make_float4(r[3] * inv, r[4] * inv, r[5] * inv, __int_as_float((int)i));
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) *count = (unsigned)n;
}
__global__ void k_rays_to_queue64(const double *__restrict__ rays, long long n, double4 *qo, double4 *qd, unsigned *count) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double *r = rays + 6 * i;
        qo[i] = make_double4(r[0], r[1], r[2], 1.0);
        qd[i] = make_double4(r[3], r[4], r[5], (double)i);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) *count = (unsigned)n;
}
__global__ void k_cvt_d2f(const double *__restrict__ src, float *__restrict__ dst, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) dst[i] = (float)src[i];
}

static int check_scene(const ptb_scene *s, void *stream = nullptr) {
    if (!s) { set_error("null scene handle"); return PTB_ERR_ARG; }
    int dev = -1;
    cudaGetDevice(&dev);
    if (dev != s->device) { set_error("scene was created on device %d, current device is %d", s->device, dev); return PTB_ERR_ARG; }
    // remember the stream: ptb_scene_destroy waits for the streams that used the scene, not for the device
    const cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(s->mu);
    if (std::find(s->streams.begin(), s->streams.end(), st) == s->streams.end()) s->streams.push_back(st);
    return PTB_OK;
}
static int check_pixels(const ptb_pixels *px) {
    if (!px || px->width <= 0 || px->height <= 0) { set_error("bad pixel set: width/height"); return PTB_ERR_ARG; }
    if ((int64_t)px->width * px->height > 0x7fffffffll) { set_error("frame too large"); return PTB_ERR_ARG; }
    if (!px->d_pixels && (px->n_shards > 1) && (px->shard < 0 || px->shard >= px->n_shards || px->tile <= 0)) {
        set_error("bad pixel set: shard %d of %d, tile %d", px->shard, px->n_shards, px->tile);
        return PTB_ERR_ARG;
    }
    return PTB_OK;
}

static size_t smem_tris(int n) { return (size_t)n * sizeof(TriRec); }
// q.work_slot rotates through the 32 work counters of a queue set; q.count_nodes: instrumented (untimed) run that counts
// nodes visited / triangles tested.  Both live in the per-call queue set: calls on other streams / threads have their own.

// L2 persistence window over a node array for the launches that follow on `st` (PTB_L2_PERSIST_MB > 0): node fetches that
// miss L1 then hit L2 instead of going to HBM whatever the ray state streaming through does to the rest of the cache.
static void l2_window(cudaStream_t st, const void *base, size_t bytes) {
    DeviceState *ds = device_state();
    if (!ds->l2_persist_bytes) return;
    cudaStreamAttrValue v;
    memset(&v, 0, sizeof(v));
    const size_t win = std::min(bytes, ds->l2_window_max);
    v.accessPolicyWindow.base_ptr = const_cast<void *>(base);
    v.accessPolicyWindow.num_bytes = base ? win : 0;
    v.accessPolicyWindow.hitRatio = (base && win) ? (float)std::min(1.0, (double)ds->l2_persist_bytes / (double)win) : 0.f;
    v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    if (cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v) != cudaSuccess) cudaGetLastError();
}

// Blocks per SM in the grid of the fused per-bounce kernel (4 are resident).  The tiles of a launch are dealt out to the blocks
// with a fixed stride, so a block's share of slow tiles is luck: with more, shorter blocks the SMs that finish early pick up
// new blocks, and the blocks that run at the same time stay on neighbouring tiles of the queues (measured, Msamples/s: 4 -> 2 843,
// 8 -> 3 013, 16 -> 3 110, 32 -> 3 172, 128 -> 3 198, 512 -> 3 185, one tile per block 2 890; PTB_FUSED_GRID=n overrides).
#ifndef PTB_FUSED2_GRID
#define PTB_FUSED2_GRID 128
#endif
static int fused_grid_blocks_per_sm() {
    static const int v = [] { const char *e = getenv("PTB_FUSED_GRID"); const int n = e ? atoi(e) : 0; return n > 0 ? n : PTB_FUSED2_GRID; }();
    return v;
}

#ifndef PTB_SHADE_GRID
#define PTB_SHADE_GRID 8
#endif
static int shade_grid_blocks_per_sm() {     // k_shade32 / k_nee_finish32 of the LBVH pipeline (PTB_SHADE_GRID_ENV=n overrides)
    static const int v = [] { const char *e = getenv("PTB_SHADE_GRID_ENV"); const int n = e ? atoi(e) : 0; return n > 0 ? n : PTB_SHADE_GRID; }();
    return v;
}

// closest-hit over queue `in` for the brute-force scene
static int launch_closest32(const ptb_scene *s, Queues<float> &q, int in, const unsigned *d_count, size_t upper,
                            bool track, float *hit_t, int *hit_tri, cudaStream_t st, const unsigned *order = nullptr,
                            const int *origin_pid = nullptr) {
    ProfScope ps_(PK_CLOSEST, st);
    if (s->d.accel == PTB_ACCEL_LBVH) {
        const int grid = grid_for(upper, PTB_BVH_THREADS, 12);
        l2_window(st, s->d.nodes, sizeof(BvhNode) * (size_t)s->d.n_tri);
        if (s->bvh_dynamic && q.work) {
            unsigned *work = q.work + (q.work_slot++ & 31);
            PTB_CUDA(cudaMemsetAsync(work, 0, sizeof(unsigned), st));
            if (track)
                k_closest_bvh32_dyn<true, false><<<grid, PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale, work, order);
            else if (origin_pid && q.count_nodes)
                k_closest_bvh32_dyn<false, true, false, true><<<grid, PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale, work, order, origin_pid);
            else if (origin_pid)
                k_closest_bvh32_dyn<false, false, false, true><<<grid, PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale, work, order, origin_pid);
            else if (q.count_nodes)
                k_closest_bvh32_dyn<false, true><<<grid, PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale, work, order);
            else
                k_closest_bvh32_dyn<false, false><<<grid, PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale, work, order);
        } else if (track)
            k_closest_bvh32<true, false><<<grid, PTB_BVH_THREADS, 0, st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale);
        else if (q.count_nodes)
            k_closest_bvh32<false, true><<<grid, PTB_BVH_THREADS, 0, st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale);
        else
            k_closest_bvh32<false, false><<<grid, PTB_BVH_THREADS, 0, st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale);
        PTB_LAUNCHED();
        return PTB_OK;
    }
    const int grid = grid_for(upper, PTB_TRACE_THREADS, 8);
    if (!track && s->use_planes)
        k_closest_planes32<false><<<grid, PTB_TRACE_THREADS, plane_smem_bytes(s->planes.n_planes), st>>>(s->d, s->planes, q.qo[in], q.qd[in], d_count, hit_t, hit_tri);
    else if (track)
        k_closest_brute32<true><<<grid, PTB_TRACE_THREADS, smem_tris(s->d.n_tri), st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale);
    else
        k_closest_brute32<false><<<grid, PTB_TRACE_THREADS, smem_tris(s->d.n_tri), st>>>(s->d, q.qo[in], q.qd[in], d_count, hit_t, hit_tri, q.unc_count, q.unc_list, s->dt_scale);
    PTB_LAUNCHED();
    return PTB_OK;
}

static int launch_closest64(const ptb_scene *s, const double4 *qo, const double4 *qd, const unsigned *d_count, size_t upper,
                            double *hit_t, int *hit_tri, cudaStream_t st, const int *origin_pid = nullptr) {
    ProfScope ps_(PK_CLOSEST, st);
    if (s->d.accel == PTB_ACCEL_LBVH)
        k_closest_bvh64<<<grid_for(upper, PTB_BVH_THREADS, 8), PTB_BVH_THREADS, 0, st>>>(s->d, qo, qd, d_count, hit_t, hit_tri, origin_pid);
    else
        k_closest_brute64<<<grid_for(upper, 128, 8), 128, 0, st>>>(s->d, qo, qd, d_count, hit_t, hit_tri);
    PTB_LAUNCHED();
    return PTB_OK;
}

}  // namespace ptb

extern "C" size_t ptb_trace_primary_workspace(int64_t n_pixels, int32_t precision) {
    const size_t n = (size_t)std::max<int64_t>(n_pixels, 1);
    return precision == PTB_PREC_FP64 ? queue_bytes<double>(n, 1) : queue_bytes<float>(n, 1);
}

extern "C" int ptb_trace_primary(const ptb_scene *s, const ptb_pixels *px, int32_t precision, int32_t flags,
                                 int32_t *d_tri, float *d_t, void *d_ws, size_t ws_bytes, void *stream) {
    int rc;
    if ((rc = check_scene(s, stream)) || (rc = check_pixels(px))) return rc;
    if (!d_tri || !d_t || !d_ws) { set_error("ptb_trace_primary: null output/workspace"); return PTB_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    BatchInfo bi;
    bi.pm = make_pixel_map(px); bi.npix = bi.pm.n; bi.S = 1; bi.s_first = 0; bi.bounce = 0; bi.seed = 0; bi.s_off = nullptr;
    const size_t n = (size_t)bi.npix;
    if (n == 0) return PTB_OK;
    if (precision == PTB_PREC_FP32) {
        Queues<float> q;
        if (!queue_carve<float>(d_ws, ws_bytes, n, 1, q)) { set_error("workspace too small: need %zu bytes", queue_bytes<float>(n, 1)); return PTB_ERR_NOMEM; }
        { ProfScope ps_(PK_GEN, st);
        k_gen<float><<<grid_for(n, 256, 8), 256, 0, st>>>(s->d, bi, q.qo[0], q.qd[0], q.L, q.count, q.unc_count, 3);
        }
        PTB_LAUNCHED();
        const bool track = !(flags & PTB_FLAG_NO_FP64_RETRACE);
        if ((rc = launch_closest32(s, q, 0, q.count, n, track, d_t, d_tri, st))) return rc;
        if (track) {
            { ProfScope ps_(PK_RETRACE, st);
            k_retrace64<<<grid_for(n / 8 + 1, 128, 4), 128, 0, st>>>(s->d, bi.pm, bi.npix, q.unc_count, q.unc_list, d_t, d_tri);
            }
            PTB_LAUNCHED();
        }
    } else if (precision == PTB_PREC_FP64) {
        Queues<double> q;
        if (!queue_carve<double>(d_ws, ws_bytes, n, 1, q)) { set_error("workspace too small: need %zu bytes", queue_bytes<double>(n, 1)); return PTB_ERR_NOMEM; }
        { ProfScope ps_(PK_GEN, st);
        k_gen<double><<<grid_for(n, 256, 8), 256, 0, st>>>(s->d, bi, q.qo[0], q.qd[0], q.L, q.count, q.unc_count, 3);
        }
        PTB_LAUNCHED();
        if ((rc = launch_closest64(s, q.qo[0], q.qd[0], q.count, n, q.hit_t, d_tri, st))) return rc;
        { ProfScope ps_(PK_OTHER, st);
        k_cvt_d2f<<<grid_for(n, 256, 8), 256, 0, st>>>(q.hit_t, d_t, (long long)n);
        }
        PTB_LAUNCHED();
    } else { set_error("unknown precision %d", precision); return PTB_ERR_ARG; }
    return PTB_OK;
}

extern "C" int ptb_trace_closest(const ptb_scene *s, const void *d_rays, int64_t n, int32_t precision,
                                 int32_t *d_tri, void *d_t, void *stream) {
    int rc;
    if ((rc = check_scene(s, stream))) return rc;
    if (!d_rays || !d_tri || !d_t || n < 0) { set_error("ptb_trace_closest: bad argument"); return PTB_ERR_ARG; }
    if (n == 0) return PTB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    // small private scratch: two vec4 queues + count
    const size_t v = precision == PTB_PREC_FP64 ? sizeof(double4) : sizeof(float4);
    char *scratch = nullptr;
    PTB_CUDA(cudaMallocAsync((void **)&scratch, 2 * v * (size_t)n + 256, st));
    struct Guard { void *p; cudaStream_t st; ~Guard() { cudaFreeAsync(p, st); } } guard_ = {scratch, st};   // every way out
    unsigned *cnt = (unsigned *)(scratch + 2 * v * (size_t)n);
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

// ---------------------------------------------------------------- path tracing
namespace ptb {

template <typename T> static size_t per_path_bytes() { return queue_bytes<T>(1 << 20, 0) / (1 << 20) + 1; }

static RecPtrs make_rec(const ptb_path_records *r, int npix, int n_samples, int bounces) {
    RecPtrs p;
    memset(&p, 0, sizeof(p));
    p.npix = npix; p.n_samples = n_samples; p.bounces = bounces; p.s_off = 0;
    if (r) {
        p.tri = r->d_tri; p.contrib = r->d_contrib; p.k_after = r->d_k_after; p.lobe = r->d_lobe; p.occ = r->d_occ;
        p.leak = r->d_leak; p.contrib64 = r->d_contrib64; p.k_after64 = r->d_k_after64;
        p.any = p.tri || p.contrib || p.k_after || p.lobe || p.occ || p.leak || p.contrib64 || p.k_after64;
    }
    return p;
}

// one batch of S samples, FP32 product path
static int run_batch32(const ptb_scene *s, const ptb_render_params *p, Queues<float> &q, BatchInfo bi, RecPtrs rec,
                       float4 *accum, cudaStream_t st) {
    int rc;
    const size_t N = (size_t)bi.npix * bi.S;
    const DScene &d = s->d;
    const bool separate = (p->flags & PTB_FLAG_SEPARATE_KERNELS) != 0;
    const bool share = !separate && p->bounces > 0;     // one primary ray + hit per pixel, read by all S samples (A-Q14)
    const bool fused = share && d.accel == PTB_ACCEL_BRUTE && s->use_planes;
    { ProfScope ps_(PK_GEN, st);
    if (share)
        k_gen_pixels<<<grid_for(fused ? (size_t)bi.npix : N, 256, 8), 256, 0, st>>>(d, bi, q.qo[0], q.qd[0], q.L, fused ? 0 : 1, q.count, q.unc_count, p->bounces + 2);
    else
        k_gen<float><<<grid_for(N, 256, 8), 256, 0, st>>>(d, bi, q.qo[0], q.qd[0], q.L, q.count, q.unc_count, p->bounces + 2);
    }
    PTB_LAUNCHED();
    const size_t n_primary = share ? (size_t)bi.npix : N;
    if (fused) {        // ptb_fused.cuh: one kernel per bounce, geometry in shared memory
        const bool track = !(p->flags & PTB_FLAG_NO_FP64_RETRACE);
        if ((rc = launch_closest32(s, q, 0, q.count, n_primary, track, q.hit_t, q.hit_tri, st))) return rc;
        if (track) {
            { ProfScope ps_(PK_RETRACE, st);
            k_retrace64<<<grid_for(n_primary / 8 + 1, 128, 4), 128, 0, st>>>(d, bi.pm, bi.npix, q.unc_count, q.unc_list, q.hit_t, q.hit_tri);
            }
            PTB_LAUNCHED();
        }
        float *ht[2] = {q.hit_t, q.hit_t2}; int *hr[2] = {q.hit_tri, q.hit_tri2};
        const size_t smem = plane_smem_bytes(s->planes.n_planes);
        for (int b = 0; b < p->bounces; ++b) {
            const int in = b & 1, out = in ^ 1;
            bi.bounce = b;
            const int last = b == p->bounces - 1;
            ProfScope ps_(PK_BOUNCE, st);
            const int grid_b = grid_for(N, PTB_SHADE_THREADS, fused_grid_blocks_per_sm());
            if (s->fused_x2) {      // packed FP32 (ptb_fused2.cuh): same results bit for bit, fewer issue slots
                const float4 *qi = b == 0 ? q.qo[0] : q.qo[in], *di = b == 0 ? q.qd[0] : q.qd[in];
                const unsigned *cin = q.count + b;
#define PTB_X2_ARGS d, s->planes, bi, qi, di, ht[in], hr[in], cin, q.count + b + 1, q.qo[out], q.qd[out], ht[out], hr[out], q.L, last, rec
                if (b == 0 && !last) k_bounce_planes32x2<true, true><<<grid_b, PTB_SHADE_THREADS, smem, st>>>(PTB_X2_ARGS);
                else if (b == 0) k_bounce_planes32x2<true, false><<<grid_b, PTB_SHADE_THREADS, smem, st>>>(PTB_X2_ARGS);
                else if (!last) k_bounce_planes32x2<false, true><<<grid_b, PTB_SHADE_THREADS, smem, st>>>(PTB_X2_ARGS);
                else k_bounce_planes32x2<false, false><<<grid_b, PTB_SHADE_THREADS, smem, st>>>(PTB_X2_ARGS);
#undef PTB_X2_ARGS
            } else if (b == 0)
                k_bounce_planes32<true><<<grid_for(N, PTB_SHADE_THREADS, 8), PTB_SHADE_THREADS, smem, st>>>(
                    d, s->planes, bi, q.qo[0], q.qd[0], ht[0], hr[0], q.count, q.count + 1, q.qo[1], q.qd[1], ht[1], hr[1], q.L, last, rec);
            else
                k_bounce_planes32<false><<<grid_for(N, PTB_SHADE_THREADS, 8), PTB_SHADE_THREADS, smem, st>>>(
                    d, s->planes, bi, q.qo[in], q.qd[in], ht[in], hr[in], q.count + b, q.count + b + 1, q.qo[out], q.qd[out], ht[out], hr[out], q.L, last, rec);
            PTB_LAUNCHED();
        }
        { ProfScope ps_(PK_REDUCE, st);
        k_reduce<float><<<grid_for((size_t)bi.npix, 256, 8), 256, 0, st>>>(q.L, accum, bi.npix, bi.S);
        }
        PTB_LAUNCHED();
        return PTB_OK;
    }
    const bool sorting = d.accel == PTB_ACCEL_LBVH && s->bvh_dynamic && s->sort_rays && q.sort.keys[0];
    const unsigned *order = nullptr;       // visiting order of the live slots of queue `in` (null: slot order)
    // plane id of the triangle every live path stands on, per output slot of the shade kernel (the second hit-record set
    // belongs to the fused small-scene pipeline and is free here): read by the shadow rays of this bounce and by the
    // closest-hit rays of the next one (ptb_bvh_dyn.cuh, PID)
    int *origin_pid = (d.accel == PTB_ACCEL_LBVH && s->bvh_dynamic && d.tri_pid) ? q.hit_tri2 : nullptr;
    for (int b = 0; b < p->bounces; ++b) {
        const int in = b & 1, out = in ^ 1;
        bi.bounce = b;
        const bool track = (b == 0) && !(p->flags & PTB_FLAG_NO_FP64_RETRACE);
        if ((rc = launch_closest32(s, q, in, q.count + b, b == 0 ? n_primary : N, track, q.hit_t, q.hit_tri, st, order,
                                   b == 0 ? nullptr : origin_pid))) return rc;
        if (track) {
            { ProfScope ps_(PK_RETRACE, st);
            k_retrace64<<<grid_for(n_primary / 8 + 1, 128, 4), 128, 0, st>>>(d, bi.pm, bi.npix, q.unc_count, q.unc_list, q.hit_t, q.hit_tri);
            }
            PTB_LAUNCHED();
        }
        { ProfScope ps_(PK_SHADE, st);
        k_shade32<<<grid_for(N, PTB_SHADE_THREADS, shade_grid_blocks_per_sm()), PTB_SHADE_THREADS, 0, st>>>(
            d, bi, q.qo[in], q.qd[in], q.hit_t, q.hit_tri, q.count + b, q.count + b + 1, q.qo[out], q.qd[out],
            q.sh_p, q.sh_d, q.sh_aux, q.L, q.cap, rec, sorting ? q.sort.keys[0] : nullptr, (share && b == 0) ? 1 : 0, origin_pid);
        }
        PTB_LAUNCHED();
        if (sorting) {      // order the survivors by the Morton key of their new origin (ptb_sort.cuh)
            ProfScope ps_(PK_SORT, st);
            const SortBufs &sb = q.sort;
            int cur = 0;
            for (int pass = 0; pass < PS_PASSES; ++pass) {
                ps_sort_pass(sb.keys[cur], pass ? sb.vals[cur] : nullptr, q.count + b + 1, 8 * pass, sb.tiles_max, sb.ghist, sb.totals,
                             pass < PS_PASSES - 1 ? sb.keys[cur ^ 1] : nullptr, sb.vals[cur ^ 1], st);
                g_launches += 3;
                cur ^= 1;
            }
            order = sb.vals[PS_PASSES & 1];     // buffers alternate, pass 0 writes vals[1]
            PTB_LAUNCHED(); --g_launches;
        }
        { ProfScope ps_(PK_SHADOW, st);
        if (d.accel == PTB_ACCEL_LBVH && s->bvh_dynamic) {
            unsigned *work = q.work + (q.work_slot++ & 31);
            PTB_CUDA(cudaMemsetAsync(work, 0, sizeof(unsigned), st));
            l2_window(st, d.nodes_occ ? d.nodes_occ : d.nodes, sizeof(BvhNode) * (size_t)(d.nodes_occ ? d.n_tri_obj : d.n_tri));
            if (origin_pid && q.count_nodes)
                k_shadow_bvh32_dyn<true, false, true><<<grid_for(3 * N, PTB_BVH_THREADS, 12), PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(d, q.sh_p, q.sh_d, q.count + b + 1, q.sh_res, q.cap, work, order, origin_pid);
            else if (origin_pid)
                k_shadow_bvh32_dyn<false, false, true><<<grid_for(3 * N, PTB_BVH_THREADS, 12), PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(d, q.sh_p, q.sh_d, q.count + b + 1, q.sh_res, q.cap, work, order, origin_pid);
            else if (q.count_nodes)
                k_shadow_bvh32_dyn<true><<<grid_for(3 * N, PTB_BVH_THREADS, 12), PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(d, q.sh_p, q.sh_d, q.count + b + 1, q.sh_res, q.cap, work, order);
            else
                k_shadow_bvh32_dyn<false><<<grid_for(3 * N, PTB_BVH_THREADS, 12), PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(d, q.sh_p, q.sh_d, q.count + b + 1, q.sh_res, q.cap, work, order);
            PTB_LAUNCHED();
            k_nee_finish32<<<grid_for(N, 256, shade_grid_blocks_per_sm()), 256, 0, st>>>(d, q.sh_p, q.sh_d, q.sh_aux, q.sh_res, q.count + b + 1, q.L, q.cap, b, rec);
        } else if (d.accel == PTB_ACCEL_LBVH) {
            if (q.count_nodes)
                k_shadow_bvh32<true><<<grid_for(N, PTB_BVH_THREADS, 12), PTB_BVH_THREADS, 0, st>>>(d, q.sh_p, q.sh_d, q.sh_aux, q.count + b + 1, q.L, q.cap, b, rec);
            else
                k_shadow_bvh32<false><<<grid_for(N, PTB_BVH_THREADS, 12), PTB_BVH_THREADS, 0, st>>>(d, q.sh_p, q.sh_d, q.sh_aux, q.count + b + 1, q.L, q.cap, b, rec);
        } else if (s->use_planes)
            k_shadow_planes32<<<grid_for(N, PTB_TRACE_THREADS, 8), PTB_TRACE_THREADS, plane_smem_bytes(s->planes.n_planes), st>>>(
                d, s->planes, q.sh_p, q.sh_d, q.sh_aux, q.count + b + 1, q.L, q.cap, b, rec);
        else
            k_shadow_brute32<<<grid_for(N, PTB_TRACE_THREADS, 8), PTB_TRACE_THREADS, smem_tris(d.n_tri_obj), st>>>(
                d, q.sh_p, q.sh_d, q.sh_aux, q.count + b + 1, q.L, q.cap, b, rec);
        }
        PTB_LAUNCHED();
    }
    { ProfScope ps_(PK_REDUCE, st);
    k_reduce<float><<<grid_for((size_t)bi.npix, 256, 8), 256, 0, st>>>(q.L, accum, bi.npix, bi.S);
    }
    PTB_LAUNCHED();
    return PTB_OK;
}

// one batch, physical mode (ptb_phys.cuh): same wavefront, ray semantics, RGB throughput, Russian roulette
static int run_batch_phys32(const ptb_scene *s, const ptb_render_params *p, Queues<float> &q, BatchInfo bi, float4 *accum,
                            cudaStream_t st) {
    const size_t N = (size_t)bi.npix * bi.S;
    const DScene &d = s->d;
    const bool lbvh = d.accel == PTB_ACCEL_LBVH;
    const int no_rr = (p->flags & PTB_FLAG_NO_ROULETTE) ? 1 : 0;
    { ProfScope ps_(PK_GEN, st);
    k_gen_pixels<<<grid_for(N, 256, 8), 256, 0, st>>>(d, bi, q.qo[0], q.qd[0], q.L, 1, q.count, q.unc_count, p->bounces + 2);
    }
    PTB_LAUNCHED();
    PTB_CUDA(cudaMemsetAsync(q.count_sh, 0, (size_t)(p->bounces + 2) * sizeof(unsigned), st));
    const bool sorting = lbvh && s->sort_rays;
    const unsigned *order = nullptr;
    const size_t smem = plane_smem_bytes(s->planes.n_planes);
    for (int b = 0; b < p->bounces; ++b) {
        const int in = b & 1, out = in ^ 1;
        bi.bounce = b;
        const size_t upper = b == 0 ? (size_t)bi.npix : N;
        { ProfScope ps_(PK_CLOSEST, st);
        if (lbvh) {
            unsigned *work = q.work + (q.work_slot++ & 31);
            PTB_CUDA(cudaMemsetAsync(work, 0, sizeof(unsigned), st));
            k_closest_bvh32_dyn<false, false, true><<<grid_for(upper, PTB_BVH_THREADS, 12), PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(
                d, q.qo[in], q.qd[in], q.count + b, q.hit_t, q.hit_tri, q.unc_count, q.unc_list, s->dt_scale, work, order);
        } else {
            k_closest_planes32<true><<<grid_for(upper, PTB_TRACE_THREADS, 8), PTB_TRACE_THREADS, smem, st>>>(
                d, s->planes, q.qo[in], q.qd[in], q.count + b, q.hit_t, q.hit_tri);
        }
        }
        PTB_LAUNCHED();
        { ProfScope ps_(PK_SHADE, st);
        k_shade_phys32<<<grid_for(N, PTB_SHADE_THREADS, 8), PTB_SHADE_THREADS, 0, st>>>(
            d, bi, q.qo[in], q.qd[in], q.kgb[in], q.hit_t, q.hit_tri, q.count + b, q.count + b + 1, q.count_sh + b, q.qo[out], q.qd[out],
            q.kgb[out], q.sh_p, q.sh_d, q.sh_aux, q.sh_w, q.sh_res, q.L, q.cap, sorting ? q.sort.keys[0] : nullptr, b == 0 ? 1 : 0,
            b == p->bounces - 1 ? 1 : 0, no_rr);
        }
        PTB_LAUNCHED();
        if (sorting && b + 1 < p->bounces) {
            ProfScope ps_(PK_SORT, st);
            const SortBufs &sb = q.sort;
            int cur = 0;
            for (int pass = 0; pass < PS_PASSES; ++pass) {
                ps_sort_pass(sb.keys[cur], pass ? sb.vals[cur] : nullptr, q.count + b + 1, 8 * pass, sb.tiles_max, sb.ghist, sb.totals,
                             pass < PS_PASSES - 1 ? sb.keys[cur ^ 1] : nullptr, sb.vals[cur ^ 1], st);
                g_launches += 3;
                cur ^= 1;
            }
            order = sb.vals[PS_PASSES & 1];
            PTB_LAUNCHED(); --g_launches;
        }
        { ProfScope ps_(PK_SHADOW, st);
        if (lbvh) {
            unsigned *work = q.work + (q.work_slot++ & 31);
            PTB_CUDA(cudaMemsetAsync(work, 0, sizeof(unsigned), st));
            k_shadow_bvh32_dyn<false, true><<<grid_for(3 * N, PTB_BVH_THREADS, 12), PTB_BVH_THREADS, PTB_DYN_SMEM, st>>>(
                d, q.sh_p, q.sh_d, q.count_sh + b, q.sh_res, q.cap, work, nullptr);
        } else {
            k_shadow_planes_phys32<<<grid_for(N, PTB_TRACE_THREADS, 8), PTB_TRACE_THREADS, smem, st>>>(
                s->planes, q.sh_p, q.sh_d, q.count_sh + b, q.sh_res, q.cap);
        }
        PTB_LAUNCHED();
        k_nee_finish_phys32<<<grid_for(N, 256, 8), 256, 0, st>>>(d, q.sh_p, q.sh_aux, q.sh_w, q.sh_res, q.count_sh + b, q.L, q.cap);
        }
        PTB_LAUNCHED();
    }
    { ProfScope ps_(PK_REDUCE, st);
    k_reduce<float><<<grid_for((size_t)bi.npix, 256, 8), 256, 0, st>>>(q.L, accum, bi.npix, bi.S);
    }
    PTB_LAUNCHED();
    return PTB_OK;
}

// one batch, FP64 validation path (reference operation order)
static int run_batch64(const ptb_scene *s, const ptb_render_params *p, Queues<double> &q, BatchInfo bi, RecPtrs rec,
                       float4 *accum, cudaStream_t st) {
    const size_t N = (size_t)bi.npix * bi.S;
    const DScene &d = s->d;
    { ProfScope ps_(PK_GEN, st);
    k_gen<double><<<grid_for(N, 256, 8), 256, 0, st>>>(d, bi, q.qo[0], q.qd[0], q.L, q.count, q.unc_count, p->bounces + 2);
    }
    PTB_LAUNCHED();
    // the plane every live path stands on (same rule as the FP32 pipeline; ptb_bvh_kernels.cuh)
    int *origin_pid = (d.accel == PTB_ACCEL_LBVH && d.tri_pid) ? q.hit_tri2 : nullptr;
    for (int b = 0; b < p->bounces; ++b) {
        const int in = b & 1, out = in ^ 1;
        bi.bounce = b;
        { int rc2 = launch_closest64(s, q.qo[in], q.qd[in], q.count + b, N, q.hit_t, q.hit_tri, st, b == 0 ? nullptr : origin_pid); if (rc2) return rc2; }
        { ProfScope ps_(PK_SHADE, st);
        k_shade64<<<grid_for(N, PTB_SHADE_THREADS, 4), PTB_SHADE_THREADS, 0, st>>>(
            d, bi, q.qo[in], q.qd[in], q.hit_tri, q.count + b, q.count + b + 1, q.qo[out], q.qd[out], q.sh_p, q.sh_d,
            q.sh_aux, q.L, q.cap, rec, origin_pid);
        }
        PTB_LAUNCHED();
        { ProfScope ps_(PK_SHADOW, st);
        if (d.accel == PTB_ACCEL_LBVH)
            k_shadow_bvh64<<<grid_for(N, PTB_BVH_THREADS, 8), PTB_BVH_THREADS, 0, st>>>(d, q.sh_p, q.sh_d, q.sh_aux, q.count + b + 1, q.L, q.cap, b, rec);
        else
            k_shadow_brute64<<<grid_for(N, 128, 8), 128, 0, st>>>(d, q.sh_p, q.sh_d, q.sh_aux, q.count + b + 1, q.L, q.cap, b, rec);
        }
        PTB_LAUNCHED();
    }
    { ProfScope ps_(PK_REDUCE, st);
    k_reduce<double><<<grid_for((size_t)bi.npix, 256, 8), 256, 0, st>>>(q.L, accum, bi.npix, bi.S);
    }
    PTB_LAUNCHED();
    return PTB_OK;
}

}  // namespace ptb

extern "C" size_t ptb_trace_paths_workspace(const ptb_render_params *p, int32_t samples_per_batch) {
    if (!p) return 0;
    const size_t npix = (size_t)pixel_count(&p->px);
    const size_t cap = npix * (size_t)std::max(1, samples_per_batch);
    return p->precision == PTB_PREC_FP64 ? queue_bytes<double>(cap, p->bounces) : queue_bytes<float>(cap, p->bounces);
}

static int trace_paths_impl(const ptb_scene *s, const ptb_render_params *p, void *d_ws, size_t ws_bytes,
                            float *d_accum, const ptb_path_records *records, const int32_t *d_sample_offset, void *stream) {
    int rc;
    if ((rc = check_scene(s, stream))) return rc;
    if (!p || (rc = check_pixels(&p->px))) return rc ? rc : PTB_ERR_ARG;
    if (!d_ws || !d_accum) { set_error("ptb_trace_paths: null workspace/accum"); return PTB_ERR_ARG; }
    if (p->sample_end < p->sample_begin || p->bounces < 0) { set_error("bad sample range / bounces"); return PTB_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    const int npix = (int)pixel_count(&p->px);
    const int n_samples = p->sample_end - p->sample_begin;
    if (npix == 0 || n_samples == 0) return PTB_OK;
    const bool f64 = p->precision == PTB_PREC_FP64;
    if (!f64 && p->precision != PTB_PREC_FP32) { set_error("unknown precision %d", p->precision); return PTB_ERR_ARG; }
    // samples per batch: as many as the workspace holds (or the caller's cap)
    int S = p->samples_per_batch > 0 ? p->samples_per_batch : n_samples;
    S = std::min(S, n_samples);
    while (S >= 1) {
        size_t need = f64 ? queue_bytes<double>((size_t)npix * S, p->bounces) : queue_bytes<float>((size_t)npix * S, p->bounces);
        if (need <= ws_bytes && (size_t)npix * S < 0x7fffffffull) break;
        S = (p->samples_per_batch > 0) ? 0 : S / 2;
    }
    if (S < 1) {
        set_error("workspace too small: %zu bytes cannot hold one sample of %d pixels (need %zu)", ws_bytes, npix,
                  f64 ? queue_bytes<double>((size_t)npix, p->bounces) : queue_bytes<float>((size_t)npix, p->bounces));
        return PTB_ERR_NOMEM;
    }
    RecPtrs rec = make_rec(records, npix, n_samples, p->bounces);
    if (p->flags & PTB_FLAG_PHYSICAL) {
        if (f64 || rec.any) { set_error("physical mode runs in FP32 and has no reference records"); return PTB_ERR_UNSUPPORTED; }
        if (s->d.accel == PTB_ACCEL_BRUTE && !s->use_planes) { set_error("physical mode needs the plane-grouped records"); return PTB_ERR_UNSUPPORTED; }
    }
    if (rec.any) {
        const size_t n = (size_t)npix * n_samples * p->bounces;
        { ProfScope ps_(PK_OTHER, st);
        k_init_records<<<grid_for(n, 256, 8), 256, 0, st>>>(rec, n);
        }
        PTB_LAUNCHED();
    }
    BatchInfo bi;
    bi.pm = make_pixel_map(&p->px); bi.npix = npix; bi.seed = p->seed; bi.bounce = 0; bi.s_off = d_sample_offset;
    Queues<float> q32; Queues<double> q64;
    if (f64) queue_carve<double>(d_ws, ws_bytes, (size_t)npix * S, p->bounces, q64);
    else queue_carve<float>(d_ws, ws_bytes, (size_t)npix * S, p->bounces, q32);
    q32.count_nodes = q64.count_nodes = (p->flags & PTB_FLAG_COUNT_NODES) != 0;
    for (int s0 = p->sample_begin; s0 < p->sample_end; s0 += S) {
        bi.S = std::min(S, p->sample_end - s0);
        bi.s_first = s0;
        rec.s_off = s0 - p->sample_begin;
        if (p->flags & PTB_FLAG_PHYSICAL) rc = run_batch_phys32(s, p, q32, bi, (float4 *)d_accum, st);
        else rc = f64 ? run_batch64(s, p, q64, bi, rec, (float4 *)d_accum, st) : run_batch32(s, p, q32, bi, rec, (float4 *)d_accum, st);
        if (rc) return rc;
    }
    return PTB_OK;
}

extern "C" int ptb_trace_paths(const ptb_scene *s, const ptb_render_params *p, void *d_ws, size_t ws_bytes,
                               float *d_accum, const ptb_path_records *records, void *stream) {
    return trace_paths_impl(s, p, d_ws, ws_bytes, d_accum, records, nullptr, stream);
}
extern "C" int ptb_trace_paths_at(const ptb_scene *s, const ptb_render_params *p, void *d_ws, size_t ws_bytes,
                                  float *d_accum, const ptb_path_records *records, const int32_t *d_sample_offset, void *stream) {
    if (!d_sample_offset) { set_error("ptb_trace_paths_at: null sample offset"); return PTB_ERR_ARG; }
    return trace_paths_impl(s, p, d_ws, ws_bytes, d_accum, records, d_sample_offset, stream);
}

extern "C" int ptb_tonemap(const float *d_accum, int32_t width, int32_t height, uint8_t *d_rgb, void *d_ws, void *stream) {
    if (!d_accum || !d_rgb || !d_ws || width <= 0 || height <= 0) { set_error("ptb_tonemap: bad argument"); return PTB_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long *mm = (unsigned long long *)d_ws;
    const int n = width * height;
    { ProfScope ps_(PK_TONEMAP, st);
    k_minmax_init<<<1, 1, 0, st>>>(mm);
    }
    PTB_LAUNCHED();
    { ProfScope ps_(PK_TONEMAP, st);
    k_minmax<<<grid_for((size_t)n, 256, 4), 256, 0, st>>>((const float4 *)d_accum, n, mm);
    }
    PTB_LAUNCHED();
    { ProfScope ps_(PK_TONEMAP, st);
    k_quantise<<<grid_for((size_t)n, 256, 8), 256, 0, st>>>((const float4 *)d_accum, width, height, mm, d_rgb);
    }
    PTB_LAUNCHED();
    return PTB_OK;
}

extern "C" int ptb_tonemap_scene(const float *d_accum, const int32_t *d_primary_tri, int32_t width, int32_t height,
                                 double tonemapping, const double *background, uint8_t *d_rgb, void *stream) {
    if (!d_accum || !d_rgb || width <= 0 || height <= 0 || !(tonemapping >= 0.0)) { set_error("ptb_tonemap_scene: bad argument"); return PTB_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    const int n = width * height;
    const float bg[3] = {background ? (float)background[0] : 0.f, background ? (float)background[1] : 0.f, background ? (float)background[2] : 0.f};
    { ProfScope ps_(PK_TONEMAP, st);
    k_quantise_scene<<<grid_for((size_t)n, 256, 8), 256, 0, st>>>((const float4 *)d_accum, d_primary_tri, width, height,
                                                                   (float)tonemapping, bg[0], bg[1], bg[2], d_rgb);
    }
    PTB_LAUNCHED();
    return PTB_OK;
}

extern "C" int ptb_scatter_accum(const ptb_pixels *px, const float *d_src, float *d_dst, void *stream) {
    int rc;
    if ((rc = check_pixels(px))) return rc;
    if (!d_src || !d_dst) { set_error("ptb_scatter_accum: null pointer"); return PTB_ERR_ARG; }
    PixelMap pm = make_pixel_map(px);
    if (pm.n == 0) return PTB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    { ProfScope ps_(PK_OTHER, st);
    k_scatter_accum<<<grid_for((size_t)pm.n, 256, 8), 256, 0, (cudaStream_t)stream>>>(pm, (const float4 *)d_src, (float4 *)d_dst);
    }
    PTB_LAUNCHED();
    return PTB_OK;
}

extern "C" int ptb_counters_read(const ptb_scene *s, ptb_counters *out, void *stream) {
    int rc;
    if ((rc = check_scene(s, stream))) return rc;
    if (!out) { set_error("null counters"); return PTB_ERR_ARG; }
    unsigned long long h[CTR_COUNT];
    PTB_CUDA(cudaMemcpyAsync(h, s->d.counters, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PTB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    out->closest_rays = h[CTR_CLOSEST]; out->shadow_rays = h[CTR_SHADOW]; out->shaded = h[CTR_SHADED];
    out->samples = h[CTR_SAMPLES]; out->retraced = h[CTR_RETRACED]; out->launches = g_launches;
    out->nodes_visited = h[CTR_NODES]; out->tris_tested = h[CTR_TRIS];
    out->nodes_visited_shadow = h[CTR_NODES_SH]; out->tris_tested_shadow = h[CTR_TRIS_SH];
    return PTB_OK;
}

extern "C" int ptb_counters_reset(const ptb_scene *s, void *stream) {
    int rc;
    if ((rc = check_scene(s, stream))) return rc;
    PTB_CUDA(cudaMemsetAsync(s->d.counters, 0, CTR_COUNT * sizeof(unsigned long long), (cudaStream_t)stream));
    g_launches = 0;
    return PTB_OK;
}

// ---------------------------------------------------------------- FP32 FMA peak microbenchmark
namespace ptb {
__global__ void __launch_bounds__(256) k_fma_peak(float *out, float a, float b, int iters) {
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
        x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
    if (x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 == 12345.678f) out[0] = x0;
}
}  // namespace ptb

extern "C" int ptb_fp32_peak(double *tflops, void *stream) {
    if (!tflops) { set_error("ptb_fp32_peak: null"); return PTB_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    float *out = nullptr;
    PTB_CUDA(cudaMalloc(&out, 16));
    const int grid = num_sms() * 8, iters = 1 << 15;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(a, st);
        k_fma_peak<<<grid, 256, 0, st>>>(out, 0.999f, 0.001f, iters);
        cudaEventRecord(b, st);
        PTB_CUDA(cudaEventSynchronize(b));
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        const double fl = 2.0 * 8.0 * iters * 256.0 * grid;
        if (rep > 0) best = std::max(best, fl / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    cudaFree(out);
    PTB_LAUNCHED();
    *tflops = best;
    return PTB_OK;
}

// ---------------------------------------------------------------- 64 B gather peak microbenchmark
// The LBVH kernels are bound by scattered 64 B fetches (two LDG.E.256 per node or triangle record), not by DRAM bytes:
// at 10^6 triangles the trees sit in L2 and re-ordered rays share L1 lines.  This measures what the memory system
// delivers for exactly that access: every lane chases its own chain of 64 B records (the next index comes out of the
// record just loaded, as a child link does) through a buffer of `bytes`, at the traversal kernels' launch shape
// (128 threads, 10 blocks per SM).  64 KB stays in L1, 64 MB in L2, > 1 GB goes to HBM.
namespace ptb {
__global__ void k_gather_fill(uint4 *buf, unsigned n) {
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint4 h = philox4x32_10(make_uint4(i, 0x9e3779b9u, 0u, 0u), 0x1234567u, 0u);
        // record = 4 x uint4; the link (next record) sits where BvhNode keeps child 0
        buf[4 * (size_t)i + 0] = make_uint4(h.y, h.z, h.w, i);
        buf[4 * (size_t)i + 1] = make_uint4(i, h.y, h.z, h.w);
        buf[4 * (size_t)i + 2] = make_uint4(h.w, i, h.y, h.z);
        buf[4 * (size_t)i + 3] = make_uint4(h.x % n, h.w, i, h.y);
    }
}
__global__ void __launch_bounds__(PTB_BVH_THREADS, PTB_BVH_MIN_BLOCKS)
k_gather_chase(const uint4 *__restrict__ buf, unsigned n, int iters, unsigned *out) {
    unsigned i = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u % n;
    float acc = 0.f;
    for (int k = 0; k < iters; ++k) {
        float4 a, b, c, d;
        ld256(buf + 4 * (size_t)i, a, b);
        ld256(buf + 4 * (size_t)i + 2, c, d);
        acc += a.x + b.y + c.z;                 // every part of the record is consumed
        i = (unsigned)__float_as_int(d.x);      // dependent: the next address comes out of this record
    }
    if (acc == 12345.678f) out[0] = i;
    if (i == 0xffffffffu) out[1] = 1;
}
}  // namespace ptb

extern "C" int ptb_gather_peak(size_t bytes, double *gfetch_per_s, void *stream) {
    if (!gfetch_per_s || bytes < 4096) { set_error("ptb_gather_peak: bad argument"); return PTB_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned n = (unsigned)std::min<size_t>(bytes / 64, 0x7fffffffu);
    uint4 *buf = nullptr; unsigned *out = nullptr;
    PTB_CUDA(cudaMalloc(&buf, (size_t)n * 64 + 64));
    struct Guard { void *a, *b; ~Guard() { cudaFree(a); if (b) cudaFree(b); } } guard_ = {buf, nullptr};
    PTB_CUDA(cudaMalloc(&out, 16));
    guard_.b = out;
    k_gather_fill<<<num_sms() * 8, 256, 0, st>>>(buf, n);
    PTB_LAUNCHED();
    const int grid = num_sms() * PTB_BVH_MIN_BLOCKS, iters = 4096;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(a, st);
        k_gather_chase<<<grid, PTB_BVH_THREADS, 0, st>>>(buf, n, iters, out);
        cudaEventRecord(b, st);
        if (cudaEventSynchronize(b) != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        if (rep > 0 && ms > 0.f) best = std::max(best, (double)grid * PTB_BVH_THREADS * iters / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    PTB_LAUNCHED();
    *gfetch_per_s = best;
    return PTB_OK;
}

// The same chase out of SHARED memory: what a block would get from the top levels of a tree staged there (north_star (4):
// "node and triangle tiles staged in shared memory").  64 B records, four LDS.128 per fetch; random record indices put
// the lanes of a quarter-warp on the same bank groups, so this is the conflict-laden rate a real walk would see.
namespace ptb {
__global__ void __launch_bounds__(PTB_BVH_THREADS, PTB_BVH_MIN_BLOCKS)
k_gather_chase_smem(const uint4 *__restrict__ buf, unsigned n, int iters, unsigned *out) {
    extern __shared__ __align__(16) uint4 s_rec[];
    for (unsigned i = threadIdx.x; i < 4 * n; i += blockDim.x) s_rec[i] = buf[i];
    __syncthreads();
    unsigned i = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u % n;
    float acc = 0.f;
    for (int k = 0; k < iters; ++k) {
        const uint4 a = s_rec[4 * i], b = s_rec[4 * i + 1], c = s_rec[4 * i + 2], d = s_rec[4 * i + 3];
        acc += __uint_as_float(a.x) + __uint_as_float(b.y) + __uint_as_float(c.z);
        i = d.x;
    }
    if (acc == 12345.678f) out[0] = i;
    if (i == 0xffffffffu) out[1] = 1;
}
}  // namespace ptb

extern "C" int ptb_gather_peak_smem(size_t bytes_per_block, double *gfetch_per_s, void *stream) {
    if (!gfetch_per_s || bytes_per_block < 1024 || bytes_per_block > 20 * 1024) { set_error("ptb_gather_peak_smem: 1 KB .. 20 KB per block (10 blocks per SM)"); return PTB_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned n = (unsigned)(bytes_per_block / 64);
    uint4 *buf = nullptr; unsigned *out = nullptr;
    PTB_CUDA(cudaMalloc(&buf, (size_t)n * 64 + 64));
    struct Guard { void *a, *b; ~Guard() { cudaFree(a); if (b) cudaFree(b); } } guard_ = {buf, nullptr};
    PTB_CUDA(cudaMalloc(&out, 16));
    guard_.b = out;
    k_gather_fill<<<num_sms(), 256, 0, st>>>(buf, n);
    PTB_LAUNCHED();
    const int grid = num_sms() * PTB_BVH_MIN_BLOCKS, iters = 4096;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(a, st);
        k_gather_chase_smem<<<grid, PTB_BVH_THREADS, (size_t)n * 64, st>>>(buf, n, iters, out);
        cudaEventRecord(b, st);
        if (cudaEventSynchronize(b) != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        if (rep > 0 && ms > 0.f) best = std::max(best, (double)grid * PTB_BVH_THREADS * iters / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    PTB_LAUNCHED();
    *gfetch_per_s = best;
    return PTB_OK;
}

// ---------------------------------------------------------------- known-answer entry (tier 1)
namespace ptb {
__global__ void k_kat(DScene sc, int what, int precision, const double *__restrict__ in, long long n, double *__restrict__ out,
                      int W, int H, uint32_t seed) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        if (what == 0) {                // rotate(axis (0,1,0), angle, v)   main.py:148-162
            const double ang = in[4 * i];
            const d3 v = mk3(in[4 * i + 1], in[4 * i + 2], in[4 * i + 3]);
            if (precision == PTB_PREC_FP64) {
                const d3 w = rotate_y64(ang, v);
                out[3 * i] = w.x; out[3 * i + 1] = w.y; out[3 * i + 2] = w.z;
            } else {                    // the product derives cos / sin from the normal (bounce32); here from the angle
                const f3 w = rotate_y32((float)cos(ang), (float)sin(ang), mk3((float)v.x, (float)v.y, (float)v.z));
                out[3 * i] = w.x; out[3 * i + 1] = w.y; out[3 * i + 2] = w.z;
            }
        } else if (what == 1) {         // pick_random_triangle + sample_random_pt   utils.py:21-46
            const uint4 w = rng4(seed, (uint32_t)in[4 * i], (uint32_t)in[4 * i + 1], (uint32_t)in[4 * i + 2], (uint32_t)in[4 * i + 3]);
            int li = -1;
            if (precision == PTB_PREC_FP64) {
                const d3 p = light_point64(sc, w, li);
                out[4 * i + 1] = p.x; out[4 * i + 2] = p.y; out[4 * i + 3] = p.z;
            } else {
                const f3 p = light_point32(sc, w, li);
                out[4 * i + 1] = p.x; out[4 * i + 2] = p.y; out[4 * i + 3] = p.z;
            }
            out[4 * i] = (double)li;
        } else {                        // make_screen_pts + make_rays   utils.py:55-69 (direction as returned: not normalised)
            d3 o, d;
            primary_ray64(sc, W, H, (int)in[i], o, d);
            out[3 * i] = d.x; out[3 * i + 1] = d.y; out[3 * i + 2] = d.z;
        }
    }
}
}  // namespace ptb

extern "C" int ptb_kat(const ptb_scene *s, int32_t what, int32_t precision, const double *d_in, int64_t n, double *d_out,
                       int32_t width, int32_t height, uint32_t seed, void *stream) {
    int rc;
    if ((rc = check_scene(s, stream))) return rc;
    if (what < 0 || what > 2 || !d_in || !d_out || n < 0 || (precision != PTB_PREC_FP32 && precision != PTB_PREC_FP64)) { set_error("ptb_kat: bad argument"); return PTB_ERR_ARG; }
    if (what == 2 && (width <= 0 || height <= 0)) { set_error("ptb_kat: width / height"); return PTB_ERR_ARG; }
    if (n == 0) return PTB_OK;
    k_kat<<<grid_for((size_t)n, 128, 4), 128, 0, (cudaStream_t)stream>>>(s->d, what, precision, d_in, (long long)n, d_out, width, height, seed);
    PTB_LAUNCHED();
    return PTB_OK;
}

extern "C" int ptb_scene_plane_culling(const ptb_scene *s) {
    return (s && s->d.accel == PTB_ACCEL_LBVH && s->d.tri_pid) ? 1 : 0;
}
