// k_draw.cu -- TfPoseEstimator.draw_humans (tf_pose/estimator.py:391-415) on a headless device buffer:
// per human, a thickness-3 circle of radius 3 at every present part (colour CocoColors[part], common.py:108-110)
// and then a thickness-3 line per CocoPairsRender pair whose two parts are present (colour CocoColors[pair order]);
// later primitives paint over earlier ones, humans in list order.
//
// Two passes, so that all primitives rasterise in parallel and the painter's order still holds:
//   k_draw_ids    one thread per primitive piece (12 polygon segments per circle, 1 per line) runs OpenCV's integer
//                 rasterisation rules (draw_raster.h) and atomicMax'es its paint-order id into a per-pixel id plane;
//   k_draw_paint  every pixel with an id takes that primitive's colour; the id plane is cleared for the next call.
#include "draw_raster.h"
#include "tfpd_internal.cuh"

namespace tfpd {
namespace {

__constant__ unsigned char c_colors[18][3] = {{255, 0, 0},   {255, 85, 0},  {255, 170, 0}, {255, 255, 0}, {170, 255, 0},
                                              {85, 255, 0},  {0, 255, 0},   {0, 255, 85},  {0, 255, 170}, {0, 255, 255},
                                              {0, 170, 255}, {0, 85, 255},  {0, 0, 255},   {85, 0, 255},  {170, 0, 255},
                                              {255, 0, 255}, {255, 0, 170}, {255, 0, 85}};
// CocoPairsRender = CocoPairs[:-2] (common.py:98-102)
__constant__ int c_render_pairs[17][2] = {{1, 2},   {1, 5},   {2, 3},  {3, 4},  {5, 6},   {6, 7},  {1, 8},  {8, 9}, {9, 10},
                                          {1, 11},  {11, 12}, {12, 13}, {1, 0},  {0, 14},  {14, 16}, {0, 15}, {15, 17}};

constexpr int kPiecesPerHuman = 18 * 12 + 17;      // 233
constexpr int kPrimsPerHuman = 18 + 17;            // paint-order slots of one human
constexpr int kAbsent = INT32_MIN;

struct EmitId {
    int *plane;
    int W;
    int id;
    __device__ void operator()(int x, int y) { atomicMax(plane + (size_t)y * W + x, id); }
};

// centers: [N][max_h][18][2] int32 (x, y) in image pixels, x == kAbsent when the part is missing
__global__ void k_draw_ids(const int *__restrict__ centers, const int *__restrict__ counts, int max_h, int H, int W,
                           int *__restrict__ ids) {
    const int n = blockIdx.y;
    const int nh = min(counts[n], max_h);
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int human = t / 256, piece = t % 256;
    if (human >= nh || piece >= kPiecesPerHuman) return;
    const int *c = centers + ((size_t)n * max_h + human) * 36;
    EmitId emit{ids + (size_t)n * H * W, W, 0};
    if (piece < 18 * 12) {
        const int part = piece / 12, seg = piece - part * 12;
        const int cx = c[2 * part], cy = c[2 * part + 1];
        if (cx == kAbsent) return;
        emit.id = human * kPrimsPerHuman + part + 1;
        tfpd_draw::cv_circle3_segment(W, H, cx, cy, seg, emit);
    } else {
        const int order = piece - 18 * 12;
        const int a = c_render_pairs[order][0], b = c_render_pairs[order][1];
        if (c[2 * a] == kAbsent || c[2 * b] == kAbsent) return;
        emit.id = human * kPrimsPerHuman + 18 + order + 1;
        tfpd_draw::cv_line(W, H, c[2 * a], c[2 * a + 1], c[2 * b], c[2 * b + 1], 3, emit);
    }
}

__global__ void k_draw_paint(unsigned char *__restrict__ img, int *__restrict__ ids, size_t npix) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
        const int id = ids[i];
        if (id > 0) {
            const int slot = (id - 1) % kPrimsPerHuman;
            const int col = slot < 18 ? slot : slot - 18;
            img[3 * i + 0] = c_colors[col][0];
            img[3 * i + 1] = c_colors[col][1];
            img[3 * i + 2] = c_colors[col][2];
            ids[i] = 0;
        }
    }
}

// centers of the humans a decode left on the device: int(x / map_w * image_w + 0.5) in double, exactly what
// estimate_paf (estimator.py:289-290) followed by draw_humans (:404) computes in Python floats
__global__ void k_centers_from_tables(const tfpd_human *__restrict__ humans, const tfpd_frame_info *__restrict__ info,
                                      int max_humans, int map_h, int map_w, int img_h, int img_w, int *__restrict__ centers,
                                      int *__restrict__ counts) {
    const int n = blockIdx.x;
    const int nh = info[n].n_humans;
    if (threadIdx.x == 0) counts[n] = nh;
    for (int t = threadIdx.x; t < nh * 18; t += blockDim.x) {
        const int human = t / 18, part = t - human * 18;
        const tfpd_human &hm = humans[(size_t)n * max_humans + human];
        int *c = centers + ((size_t)n * max_humans + human) * 36 + 2 * part;
        if (hm.cid[part] < 0) {
            c[0] = c[1] = kAbsent;
        } else {
            const double fx = (double)hm.x[part] / (double)map_w, fy = (double)hm.y[part] / (double)map_h;
            c[0] = (int)(fx * (double)img_w + 0.5);
            c[1] = (int)(fy * (double)img_h + 0.5);
        }
    }
}

}  // namespace

int launch_draw(const int *d_centers, const int *d_counts, int max_h, unsigned char *d_img, int *d_ids, int N, int H,
                int W, cudaStream_t s) {
    dim3 grid(max_h, N);                // 256 threads = the pieces of one human
    k_draw_ids<<<grid, 256, 0, s>>>(d_centers, d_counts, max_h, H, W, d_ids);
    const size_t npix = (size_t)N * H * W;
    const int blocks = (int)((npix + 255) / 256 < (size_t)(148 * 16) ? (npix + 255) / 256 : (size_t)(148 * 16));
    k_draw_paint<<<blocks, 256, 0, s>>>(d_img, d_ids, npix);
    return cudaGetLastError() == cudaSuccess ? 2 : -1;
}

int launch_centers_from_tables(const Workspace &ws, int N, int map_h, int map_w, int img_h, int img_w, int *d_centers,
                               int *d_counts, cudaStream_t s) {
    k_centers_from_tables<<<N, 128, 0, s>>>(ws.humans, ws.info, ws.max_humans, map_h, map_w, img_h, img_w, d_centers,
                                            d_counts);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
