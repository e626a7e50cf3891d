// iou.cu — dense IoU / intersection matrices with the reference's float64 arithmetic
// (libs/boxes/bbox.pyx).  These exist for API parity (callers that want the [N,K] matrix);
// the anchor-target path never materialises it (anchor_target.cu).
//
// Tiling: a block owns 256 rows x up to 64 query boxes.  Query boxes and their areas are staged
// in shared memory once per block; each thread keeps its row box in registers and streams its
// 64 outputs.  Writes are row-major [n,k]: a thread's consecutive k are contiguous, and the
// block covers a dense 256 x 64 x 8 B tile.  Compiled with --fmad=false: (a*b)+c-(d*e) must
// round exactly as the C reference does.
#include "common.cuh"

using namespace detbox;

constexpr int KT = 64;

template <int MODE>   // 0 IoU (bbox.pyx:15-55), 1 inter/area(query) (bbox.pyx:57-94)
__global__ void __launch_bounds__(256)
pairwise_kernel(const double *__restrict__ boxes, long n, const double *__restrict__ query, long k,
                double *__restrict__ out)
{
    __shared__ double q[KT][5];
    const long k0 = (long)blockIdx.y * KT;
    const int kn = (int)min((long)KT, k - k0);
    for (int i = threadIdx.x; i < kn; i += blockDim.x) {
        const double *p = query + (k0 + i) * 4;
        const double x1 = p[0], y1 = p[1], x2 = p[2], y2 = p[3];
        q[i][0] = x1; q[i][1] = y1; q[i][2] = x2; q[i][3] = y2;
        q[i][4] = (x2 - x1 + 1) * (y2 - y1 + 1);
    }
    __syncthreads();
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const double bx1 = boxes[r * 4], by1 = boxes[r * 4 + 1], bx2 = boxes[r * 4 + 2], by2 = boxes[r * 4 + 3];
    const double barea = (bx2 - bx1 + 1) * (by2 - by1 + 1);
    double *o = out + r * k + k0;
    for (int j = 0; j < kn; ++j) {
        double v = 0.0;
        const double iw = fmin(bx2, q[j][2]) - fmax(bx1, q[j][0]) + 1;
        if (iw > 0) {
            const double ih = fmin(by2, q[j][3]) - fmax(by1, q[j][1]) + 1;
            if (ih > 0) {
                if (MODE == 0) {
                    const double ua = barea + q[j][4] - iw * ih;
                    v = iw * ih / ua;
                } else {
                    v = iw * ih / q[j][4];
                }
            }
        }
        o[j] = v;
    }
}

// bbox.pyx:96-135: flag[n] = any_k( inter/area(box_n) > thr )
__global__ void __launch_bounds__(256)
exclude_ignored_kernel(const double *__restrict__ boxes, long n, const double *__restrict__ ign, long k,
                       double thr, int64_t *__restrict__ out)
{
    __shared__ double q[KT][4];
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    double bx1 = 0, by1 = 0, bx2 = 0, by2 = 0, barea = 1;
    if (r < n) {
        bx1 = boxes[r * 4]; by1 = boxes[r * 4 + 1]; bx2 = boxes[r * 4 + 2]; by2 = boxes[r * 4 + 3];
        barea = (bx2 - bx1 + 1.0) * (by2 - by1 + 1.0);
    }
    int flag = 0;
    for (long k0 = 0; k0 < k; k0 += KT) {
        const int kn = (int)min((long)KT, k - k0);
        __syncthreads();
        for (int i = threadIdx.x; i < kn * 4; i += blockDim.x) q[i >> 2][i & 3] = ign[k0 * 4 + i];
        __syncthreads();
        for (int j = 0; j < kn; ++j) {
            const double iw = fmin(bx2, q[j][2]) - fmax(bx1, q[j][0]) + 1;
            if (iw > 0) {
                const double ih = fmin(by2, q[j][3]) - fmax(by1, q[j][1]) + 1;
                if (ih > 0 && iw * ih / barea > thr) flag = 1;
            }
        }
    }
    if (r < n) out[r] = flag;
}

static int launch_pairwise(int mode, const double *b, long n, const double *q, long k, double *out, void *stream)
{
    if (n < 0 || k < 0) return DETBOX_ERR_INVALID;
    if (n == 0 || k == 0) return DETBOX_OK;
    if (!b || !q || !out) return DETBOX_ERR_INVALID;
    dim3 grid((unsigned)((n + 255) / 256), (unsigned)((k + KT - 1) / KT));
    if (mode == 0) pairwise_kernel<0><<<grid, 256, 0, as_stream(stream)>>>(b, n, q, k, out);
    else pairwise_kernel<1><<<grid, 256, 0, as_stream(stream)>>>(b, n, q, k, out);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_bbox_overlaps_f64(const double *b, long n, const double *q, long k, double *out, void *stream)
{ return launch_pairwise(0, b, n, q, k, out, stream); }

extern "C" int detbox_bbox_intersections_f64(const double *b, long n, const double *q, long k, double *out, void *stream)
{ return launch_pairwise(1, b, n, q, k, out, stream); }

extern "C" int detbox_bbox_exclude_ignored_f64(const double *b, long n, const double *ign, long k,
                                               double thr, int64_t *out, void *stream)
{
    if (n < 0 || k < 0) return DETBOX_ERR_INVALID;
    if (n == 0) return DETBOX_OK;
    if (!b || !out || (k > 0 && !ign)) return DETBOX_ERR_INVALID;
    exclude_ignored_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(b, n, ign, k, thr, out);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}
