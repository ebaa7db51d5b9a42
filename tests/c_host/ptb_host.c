/* ptb_host.c - a host program in plain C over the C ABI of include/ptb.h: no Python, no torch.
 *
 *   ptb_host SCENE.bin OUT.bin WIDTH HEIGHT SPP BOUNCES SEED
 *
 * SCENE.bin (written by tests/test_c_host.py from a flattened scene): int32 n_tri, n_tri_obj, n_obj, n_light_tri, accel,
 * pad; then doubles: verts[n_tri*9], normals[n_tri*3], mat[n_obj*8], light_areas[n_light_tri], eye[3], ortho[4],
 * ambient, light_color[3]; then int32 tri_obj[n_tri].
 * OUT.bin: int32 primary triangle ids [W*H], float accum [W*H*4], uint8 image [H*W*3].
 * The steps are the ones a maintainer of the reference would take in place of main.py:186-289:
 * ptb_scene_create -> ptb_trace_primary -> ptb_trace_paths -> ptb_tonemap, device buffers from cudaMalloc. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <cuda_runtime_api.h>
#include "ptb.h"

#define CK(call) do { int rc_ = (call); if (rc_ != 0) { fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, ptb_last_error()); return 2; } } while (0)
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_)); return 3; } } while (0)

static int read_all(FILE *f, void *dst, size_t bytes) { return fread(dst, 1, bytes, f) == bytes ? 0 : -1; }

int main(int argc, char **argv) {
    if (argc < 8) { fprintf(stderr, "usage: %s scene.bin out.bin W H spp bounces seed\n", argv[0]); return 1; }
    const int W = atoi(argv[3]), H = atoi(argv[4]), spp = atoi(argv[5]), bounces = atoi(argv[6]);
    const unsigned seed = (unsigned)strtoul(argv[7], NULL, 10);
    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 1; }
    int32_t hdr[6];
    if (read_all(f, hdr, sizeof hdr)) return 1;
    ptb_scene_desc d;
    memset(&d, 0, sizeof d);
    d.n_tri = hdr[0]; d.n_tri_obj = hdr[1]; d.n_obj = hdr[2]; d.n_light_tri = hdr[3]; d.accel = hdr[4];
    double *verts = malloc(sizeof(double) * 9 * d.n_tri), *normals = malloc(sizeof(double) * 3 * d.n_tri);
    double *mat = malloc(sizeof(double) * 8 * d.n_obj), *areas = malloc(sizeof(double) * d.n_light_tri);
    int32_t *tri_obj = malloc(sizeof(int32_t) * d.n_tri);
    double tail[11];
    if (read_all(f, verts, sizeof(double) * 9 * d.n_tri) || read_all(f, normals, sizeof(double) * 3 * d.n_tri) ||
        read_all(f, mat, sizeof(double) * 8 * d.n_obj) || read_all(f, areas, sizeof(double) * d.n_light_tri) ||
        read_all(f, tail, sizeof tail) || read_all(f, tri_obj, sizeof(int32_t) * d.n_tri)) { fprintf(stderr, "short scene file\n"); return 1; }
    fclose(f);
    d.verts = verts; d.normals = normals; d.mat = mat; d.light_areas = areas; d.tri_obj = tri_obj;
    memcpy(d.eye, tail, 3 * sizeof(double)); memcpy(d.ortho, tail + 3, 4 * sizeof(double));
    d.ambient = tail[7]; memcpy(d.light_color, tail + 8, 3 * sizeof(double));

    ptb_scene *scene = NULL;
    CK(ptb_scene_create(&d, NULL, &scene));                          /* Scene(path) hand-off, main.py:167 */

    ptb_pixels px;
    memset(&px, 0, sizeof px);
    px.width = W; px.height = H; px.n_shards = 1; px.tile = H;
    const int64_t npix = ptb_pixel_count(&px);
    int32_t *d_tri; float *d_t, *d_accum; uint8_t *d_rgb; void *d_ws;
    CU(cudaMalloc((void **)&d_tri, 4 * npix)); CU(cudaMalloc((void **)&d_t, 4 * npix));
    CU(cudaMalloc((void **)&d_accum, 16 * npix)); CU(cudaMalloc((void **)&d_rgb, 3 * npix));
    CU(cudaMemset(d_accum, 0, 16 * npix));

    size_t ws_bytes = ptb_trace_primary_workspace(npix, PTB_PREC_FP32);
    ptb_render_params p;
    memset(&p, 0, sizeof p);
    p.px = px; p.sample_begin = 0; p.sample_end = spp; p.bounces = bounces; p.seed = seed; p.precision = PTB_PREC_FP32;
    const size_t ws_paths = ptb_trace_paths_workspace(&p, spp);      /* all samples in one batch */
    if (ws_paths > ws_bytes) ws_bytes = ws_paths;
    if (ws_bytes < 256) ws_bytes = 256;
    CU(cudaMalloc(&d_ws, ws_bytes));

    CK(ptb_trace_primary(scene, &px, PTB_PREC_FP32, 0, d_tri, d_t, d_ws, ws_bytes, NULL));   /* main.py:199-204 */
    CK(ptb_trace_paths(scene, &p, d_ws, ws_bytes, d_accum, NULL, NULL));                      /* main.py:186-271 */
    CK(ptb_tonemap(d_accum, W, H, d_rgb, d_ws, NULL));                                       /* utils.py:150-161 */
    CU(cudaDeviceSynchronize());

    int32_t *tri = malloc(4 * npix); float *accum = malloc(16 * npix); uint8_t *rgb = malloc(3 * npix);
    CU(cudaMemcpy(tri, d_tri, 4 * npix, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(accum, d_accum, 16 * npix, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(rgb, d_rgb, 3 * npix, cudaMemcpyDeviceToHost));
    FILE *o = fopen(argv[2], "wb");
    if (!o) { perror(argv[2]); return 1; }
    fwrite(tri, 4, npix, o); fwrite(accum, 16, npix, o); fwrite(rgb, 3, npix, o);
    fclose(o);

    ptb_counters c;
    CK(ptb_counters_read(scene, &c, NULL));
    printf("ptb_host ok: %lld pixels, %llu closest-hit rays, %llu shadow rays, %llu kernel launches\n", (long long)npix,
           (unsigned long long)c.closest_rays, (unsigned long long)c.shadow_rays, (unsigned long long)c.launches);
    CK(ptb_scene_destroy(scene));
    cudaFree(d_tri); cudaFree(d_t); cudaFree(d_accum); cudaFree(d_rgb); cudaFree(d_ws);
    return 0;
}
