// Host side of the ORB extractor handle (snk_orb): geometry tables, device memory, launch sequence.
// Mirrors Saiga::ORBExtractor (saiga/vision/features/ORBExtractor.cpp:82-289) stage by stage; see orb_kernels.cuh.
#include "orb_kernels.cuh"
#include "image_kernels.cuh"

#include <nvtx3/nvToolsExt.h>  // header-only NVTX 3: ranges cost nothing unless a profiler is attached

#include <algorithm>
#include <cmath>
#include <vector>

#define SNK_PATTERN_QUAL static const
#include "../../include/snakeorb_pattern.h"

namespace snk
{
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int select_device(int device)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
    {
        set_error("no CUDA device: %s (libsnakeorb has no CPU fallback)", cudaGetErrorString(e));
        return SNK_ERR_CUDA;
    }
    if (device < 0 || device >= n)
    {
        set_error("device %d out of range (%d devices)", device, n);
        return SNK_ERR_INVALID;
    }
    SNK_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    SNK_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
    {
        set_error("device %d is sm_%d%d; libsnakeorb is built for sm_100a only", device, prop.major, prop.minor);
        return SNK_ERR_UNSUPPORTED;
    }
    return SNK_OK;
}

// Saiga::iRound (saiga/core/math/imath.h:73)
static inline int iround_f(float v) { return (int)(v + (v >= 0 ? 0.5f : -0.5f)); }

// cv::resize INTER_LINEAR coefficient tables (11-bit); x taps get the fraction zeroed at the borders, y taps do not
static void resize_tables(int ssize, int dsize, std::vector<int>& ofs, std::vector<short>& coef, bool clamp_coef)
{
    ofs.resize(dsize);
    coef.resize(2 * (size_t)dsize);
    const double scale = 1.0 / ((double)dsize / ssize);
    for (int d = 0; d < dsize; ++d)
    {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s   = (int)std::floor(f);
        f -= s;
        if (clamp_coef)
        {
            if (s < 0) f = 0, s = 0;
            if (s >= ssize - 1) f = 0, s = ssize - 1;
        }
        ofs[d]          = s;
        coef[2 * d]     = (short)std::lrintf((1.f - f) * 2048.f);
        coef[2 * d + 1] = (short)std::lrintf(f * 2048.f);
    }
}
}  // namespace snk

using namespace snk;

struct snk_orb
{
    snk_orb_params prm{};
    // ScalePyramid (saiga/vision/util/ScalePyramid.cpp:16-55)
    double scale[SNK_MAX_LEVELS]{}, inv_scale[SNK_MAX_LEVELS]{};
    int features[SNK_MAX_LEVELS]{};
    int max_kp = 0;

    cudaStream_t own_stream = nullptr, stream = nullptr;
    // the blur only depends on the pyramid: it runs on a side stream next to FAST + distribution and joins before the
    // descriptor kernel (serialised on the main stream while per-stage profiling is on, so stage times stay separable)
    cudaStream_t side_stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaStream_t dense_stream = nullptr;  // the dense pass of the quadtree distribution runs beside the ordinary one
    cudaEvent_t ev_fork2 = nullptr, ev_join2 = nullptr;
    cudaEvent_t ev_sync = nullptr;  // orders another stream against this handle's stream (snk_orb_order_*)
    int rows = 0, cols = 0;  // geometry the buffers are sized for (0 = none yet)
    int last_n = 0;

    PyrDev P{};
    std::vector<CellDev> cells;
    TmaMaps maps{};                 // one tensor map per level for the FAST tile loads
    int box_h[SNK_MAX_LEVELS]{};
    int fast_rows = 0, fast_list = 0;  // shared-memory sizing of k_fast: tallest TMA box, largest detection zone
    size_t fast_smem = 0;
    const void* map0_ptr = nullptr;  // what the level-0 map currently describes (it may alias a caller buffer)
    int map0_pitch = 0;
    size_t map0_stride = 0;
    std::vector<TileDev> tiles;
    std::vector<BlurTask> blur_tasks;  // k_blur_mma: (level, 24-column strip, 128-row band)
    BlurTask* d_blur_tasks = nullptr;
    // blur kernel (SNK_BLUR_MMA): default k_blur_umma (tcgen05, 0.39 ms per 512 EuRoC images); 1 = k_blur_mma (mma.sync, 0.585 ms;
    // also taken for pyramids of more than 8 levels); 0 = k_blur (dp4a, 0.62 ms)
    bool blur_mma  = true;
    bool blur_umma = true;
    UmmaBlurMaps umma_maps{};
    std::vector<UmmaBlurTask> umma_tasks;
    UmmaBlurTask* d_umma_tasks   = nullptr;
    int umma_tiles_per_img      = 0;
    int n_sm                    = 148;
    unsigned char* d_umma_bands = nullptr;
    int node_cap = 0;
    size_t qt_smem = 0;
    int qt_big_levels = 0, qt_big_cap = 0;  // levels [0, qt_big_levels) run in their own launch with qt_big_cap keypoints in smem
    size_t qt_smem_big = 0;

    uint8_t* d_rgba = nullptr;   // staging for host RGBA input (snk_orb_enqueue_rgba), grown on demand
    size_t rgba_bytes = 0;
    uint8_t* d_pyr = nullptr;    // levels 0..L-1 (level 0 unused when the caller's device buffer is aliased)
    uint8_t* d_gauss = nullptr;
    LevelDev own_l0{};           // level-0 descriptor of the internal buffer
    ResizeQuad* d_quads[SNK_MAX_LEVELS]{};
    ResizeRow* d_rows[SNK_MAX_LEVELS]{};
    // k_describe_tma (default for up to 8 levels; SNK_DESCRIBE_TMA=0 selects k_describe): per level the tensor maps of its patches
    bool describe_tma = true;
    DescMaps desc_maps{};
    // k_resize_tma (default; SNK_RESIZE_TMA=0 selects k_resize): per destination level the source's tensor map and its box
    bool resize_tma = true;
    CUtensorMap rs_map[SNK_MAX_LEVELS]{};
    int rs_box_w[SNK_MAX_LEVELS]{}, rs_box_h[SNK_MAX_LEVELS]{};
    uint2* d_moment_tab = nullptr;  // ComputeAngle disc as per-word dp4a weights (orb_kernels.cuh: DS_MOM_TABLE)
    float2* d_pattern_t = nullptr;  // rBRIEF pattern as float2, transposed to [point-in-byte][lane]
    CellDev* d_cells = nullptr;
    GroupDev* d_groups = nullptr;  // k_fast_group: runs of up to FASTG_CELLS neighbouring cells
    std::vector<GroupDev> groups;
    int fastg_list = 0;
    bool fast_chunk = false;  // SNK_FAST_CHUNK=1: compass pass of k_fast_group on 16-pixel chunks (measured 1.179 against 1.137 ms on words)
    size_t fastg_smem = 0;
    bool per_cell_fast = false;    // SNK_FAST_PER_CELL=1 selects the one-cell-at-a-time kernel (A/B comparisons)
    TileDev* d_tiles = nullptr;
    uint32_t *d_cand = nullptr, *d_workA = nullptr, *d_workB = nullptr, *d_sel = nullptr;
    int *d_cell_count = nullptr, *d_cell_off = nullptr, *d_sel_count = nullptr, *d_counts = nullptr, *d_err = nullptr;
    snk_keypoint_f32* d_kps = nullptr;
    uint64_t* d_desc = nullptr;
    int* h_counts = nullptr;  // pinned
    int* h_err = nullptr;     // pinned

    // per-stage timing: a ring of event sets so that up to PROF_RING enqueues can be in flight without a host sync
    static constexpr int PROF_RING = 32;
    bool profiling = false;
    cudaEvent_t ev[PROF_RING][SNK_ORB_NUM_STAGES]{};
    bool ev_pending[PROF_RING]{};
    int ev_launches[PROF_RING][SNK_ORB_NUM_STAGES]{};
    int ev_next = 0;
    float stage_ms[SNK_ORB_NUM_STAGES]{};
    int stage_launches[SNK_ORB_NUM_STAGES]{};

    void free_buffers()
    {
        auto F = [](auto*& p) {
            if (p) cudaFree(p);
            p = nullptr;
        };
        F(d_pyr), F(d_gauss), F(d_cells), F(d_groups), F(d_tiles), F(d_blur_tasks), F(d_umma_tasks), F(d_umma_bands), F(d_cand), F(d_workA), F(d_workB), F(d_sel), F(d_cell_count),
            F(d_cell_off), F(d_sel_count), F(d_counts), F(d_err), F(d_kps), F(d_desc);
        for (int l = 0; l < SNK_MAX_LEVELS; ++l) F(d_quads[l]), F(d_rows[l]);
        rows = cols = 0;
    }
};

// dynamic shared memory opt-in of k_fast / k_fast_group / k_distribute: one ceiling for every handle of the process
static constexpr int SMEM_CEILING = 200 * 1024;

// cuTensorMapEncodeTiled through the runtime's driver entry point (no libcuda link dependency)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static int encode_level_map(snk_orb* h, int l, const void* base, int pitch, size_t img_stride)
{
    static EncodeTiledFn fn = nullptr;
    if (!fn)
    {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        SNK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        SNK_REQUIRE(p && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
        fn = (EncodeTiledFn)p;
    }
    const LevelDev& L = h->P.lv[l];
    SNK_REQUIRE(((uintptr_t)base % 16) == 0 && pitch % 16 == 0 && img_stride % 16 == 0, "internal: TMA alignment");
    const cuuint64_t dims[3]    = {(cuuint64_t)L.w, (cuuint64_t)L.h, (cuuint64_t)h->prm.max_batch};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)img_stride};
    const cuuint32_t box[3]     = {(cuuint32_t)(h->per_cell_fast ? FAST_PITCH : FASTG_PITCH), (cuuint32_t)h->box_h[l], 1};
    const cuuint32_t estr[3]    = {1, 1, 1};
    CUresult r = fn(&h->maps.level[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    SNK_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for level %d (CUresult %d)", l, (int)r);
    return SNK_OK;
}

// tensor maps of k_describe_tma for one level: the level image as 48 x 31 boxes, its blurred copy as 64 x 39 boxes
static int encode_describe_maps(snk_orb* h, int l, const void* src, int pitch, size_t img_stride, bool with_gauss)
{
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    SNK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    SNK_REQUIRE(p && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
    const LevelDev& L = h->P.lv[l];
    SNK_REQUIRE(((uintptr_t)src % 16) == 0 && pitch % 16 == 0 && img_stride % 16 == 0, "internal: TMA alignment");
    const cuuint64_t dims[3] = {(cuuint64_t)L.w, (cuuint64_t)L.h, (cuuint64_t)h->prm.max_batch};
    const cuuint32_t estr[3] = {1, 1, 1};
    {
        const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)img_stride};
        const cuuint32_t box[3]     = {DT_IMG_W, DT_IMG_H, 1};
        CUresult r = ((EncodeTiledFn)p)(&h->desc_maps.img[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(src), dims, strides, box,
                                        estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        SNK_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for the moment patches of level %d (CUresult %d)", l, (int)r);
    }
    if (with_gauss)
    {
        const cuuint64_t strides[2] = {(cuuint64_t)L.gauss_pitch, (cuuint64_t)L.gauss_stride};
        const cuuint32_t box[3]     = {DT_G_W, DT_G_H, 1};
        CUresult r = ((EncodeTiledFn)p)(&h->desc_maps.gauss[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, L.gauss, dims, strides, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        SNK_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for the descriptor patches of level %d (CUresult %d)", l, (int)r);
    }
    return SNK_OK;
}

// tensor map of k_resize_tma for destination level l: the source level as rs_box_w x rs_box_h boxes, no swizzle
static int encode_resize_map(snk_orb* h, int l, const void* src, int pitch, size_t img_stride)
{
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    SNK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    SNK_REQUIRE(p && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
    const LevelDev& S = h->P.lv[l - 1];
    SNK_REQUIRE(((uintptr_t)src % 16) == 0 && pitch % 16 == 0 && img_stride % 16 == 0, "internal: TMA alignment");
    const cuuint64_t dims[3]    = {(cuuint64_t)S.w, (cuuint64_t)S.h, (cuuint64_t)h->prm.max_batch};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)img_stride};
    const cuuint32_t box[3]     = {(cuuint32_t)h->rs_box_w[l], (cuuint32_t)h->rs_box_h[l], 1};
    const cuuint32_t estr[3]    = {1, 1, 1};
    CUresult r = ((EncodeTiledFn)p)(&h->rs_map[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(src), dims, strides, box, estr,
                                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    SNK_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for the resize source of level %d (CUresult %d)", l, (int)r);
    return SNK_OK;
}

// tensor maps of k_blur_umma for one level: the source as 128 x 96 boxes with the 128-byte swizzle of the MMA operands,
// the destination (the level's blurred copy) as dense 96 x 80 boxes
static int encode_umma_maps(snk_orb* h, int l, const void* src, int pitch, size_t img_stride, bool with_dst)
{
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    SNK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    SNK_REQUIRE(p && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
    EncodeTiledFn fn  = (EncodeTiledFn)p;
    const LevelDev& L = h->P.lv[l];
    SNK_REQUIRE(((uintptr_t)src % 16) == 0 && pitch % 16 == 0 && img_stride % 16 == 0, "internal: TMA alignment");
    const cuuint64_t dims[3] = {(cuuint64_t)L.w, (cuuint64_t)L.h, (cuuint64_t)h->prm.max_batch};
    const cuuint32_t estr[3] = {1, 1, 1};
    // (L2 promotion 64 / 128 / 256 B measured: 0.352 / 0.355 / 0.377 ms against 0.356 ms without)
    const CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_NONE;
    {
        const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)img_stride};
        const cuuint32_t box[3]     = {UB_K0, UB_ROWS_IN, 1}, boxx[3] = {UB_K1, UB_ROWS_IN, 1};
        CUresult r = fn(&h->umma_maps.src[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(src), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        SNK_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for the blur source of level %d (CUresult %d)", l, (int)r);
        r = fn(&h->umma_maps.srcx[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(src), dims, strides, boxx, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        SNK_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for the blur source (side box) of level %d (CUresult %d)", l, (int)r);
    }
    if (with_dst)
    {
        const cuuint64_t strides[2] = {(cuuint64_t)L.gauss_pitch, (cuuint64_t)L.gauss_stride};
        const cuuint32_t box[3]     = {UB_OUT_COLS, UB_ROWS_OUT, 1};
        SNK_REQUIRE(((uintptr_t)L.gauss % 16) == 0 && L.gauss_stride % 16 == 0, "internal: TMA alignment");
        CUresult r = fn(&h->umma_maps.dst[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, L.gauss, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        SNK_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for the blur destination of level %d (CUresult %d)", l, (int)r);
    }
    return SNK_OK;
}

// cv::GaussianBlur of every level (ORBExtractor.cpp:213): the tensor-core kernel, or the dp4a one (SNK_BLUR_MMA=0) for A/B runs
static void launch_blur(snk_orb* h, const PyrDev& P, int n_imgs, cudaStream_t st)
{
    if (h->blur_umma)
    {
        // persistent: two CTAs per SM (their TMEM columns), each with an equal share of the launch's tiles
        const long long total = (long long)h->umma_tiles_per_img * n_imgs;
        static const int ctas_per_sm = [] {
            const char* e = getenv("SNK_UMMA_CTAS");
            return e && e[0] >= '1' && e[0] <= '4' ? e[0] - '0' : UB_CTAS_PER_SM;
        }();
        k_blur_umma<<<(unsigned)std::min<long long>((long long)ctas_per_sm * h->n_sm, total), UB_NT, UB_SMEM, st>>>(
            h->umma_maps, h->d_umma_tasks, (int)h->umma_tasks.size(), n_imgs, h->umma_tiles_per_img, h->d_umma_bands, h->d_err);
    }
    else if (h->blur_mma)
        k_blur_mma<<<dim3((unsigned)(h->blur_tasks.size() + BM_WARPS - 1) / BM_WARPS, n_imgs), BM_WARPS * 32, 0, st>>>(
            P, h->d_blur_tasks, (int)h->blur_tasks.size());
    else
        k_blur<<<dim3((unsigned)(h->tiles.size() + 7) / 8, n_imgs), 256, 0, st>>>(P, h->d_tiles, (int)h->tiles.size());
}

static int orb_allocate(snk_orb* h, int rows, int cols)
{
    if (h->rows == rows && h->cols == cols) return SNK_OK;
    h->free_buffers();
    {
        const char* e = getenv("SNK_FAST_PER_CELL");  // decides the TMA box width, so it is read before the maps are built
        h->per_cell_fast = e && e[0] == '1';
    }
    const int L = h->prm.nlevels, B = h->prm.max_batch;
    SNK_REQUIRE(rows <= 4000 && cols <= 4000, "image %dx%d too large (max 4000x4000)", cols, rows);
    PyrDev& P = h->P;
    memset(&P, 0, sizeof(P));
    P.nlevels = L;
    P.th_ini  = h->prm.th_fast;
    P.th_min  = h->prm.th_fast_min;

    // level sizes (ORBExtractor::AllocatePyramid, ORBExtractor.cpp:244-246) and FAST cell grid (:99-127)
    h->cells.clear();
    h->fast_rows = h->fast_list = 0;
    h->tiles.clear();
    h->blur_tasks.clear();
    {
        const char* e = getenv("SNK_BLUR_MMA");
        // default: the tcgen05 kernel (k_blur_umma; its tensor-map parameter block holds 8 levels); 1: the mma.sync kernel
        // k_blur_mma, also the fallback for deeper pyramids; 0: the dp4a kernel k_blur
        h->blur_mma   = !(e && e[0] == '0');
        h->blur_umma  = !(e && (e[0] == '0' || e[0] == '1')) && L <= 8;
        h->umma_tasks.clear();
        h->umma_tiles_per_img = 0;
    }
    size_t pyr_bytes = 0, gauss_bytes = 0;
    std::vector<size_t> pyr_off(L), gauss_off(L);
    int cell_cap = 1, sel_acc = 0;
    for (int l = 0; l < L; ++l)
    {
        LevelDev& lv   = P.lv[l];
        const float s  = (float)h->inv_scale[l];
        lv.h           = iround_f(rows * s);
        lv.w           = iround_f(cols * s);
        const int minB = 16, maxBX = lv.w - 16, maxBY = lv.h - 16;
        const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
        const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f);
        SNK_REQUIRE(nCols >= 1 && nRows >= 1,
                    "level %d is %dx%d: too small for the 30-px FAST cell grid (the reference divides by zero here); "
                    "use fewer levels",
                    l, lv.w, lv.h);
        const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
        // level 0 only ever receives the caller's bytes: a dense pitch makes the host->device copy one contiguous
        // transfer when the caller's rows are dense too
        lv.pitch       = (l == 0) ? align_up(lv.w, 16) : align_up(lv.w, 64);  // TMA: row pitch multiple of 16 bytes
        lv.gauss_pitch = align_up(lv.w, 64);
        lv.img_stride  = (size_t)lv.pitch * lv.h;
        lv.gauss_stride = (size_t)lv.gauss_pitch * lv.h;
        pyr_off[l]   = pyr_bytes;
        gauss_off[l] = gauss_bytes;
        pyr_bytes += align_up(lv.img_stride * B, (size_t)256);
        gauss_bytes += align_up(lv.gauss_stride * B, (size_t)256);
        lv.quota     = h->features[l];
        lv.scale     = (float)h->scale[l];
        lv.size      = (float)(int)(31 * h->scale[l]);
        lv.cell_base = (int)h->cells.size();
        for (int i = 0; i < nRows; ++i)
        {
            const int iniY = minB + i * hCell;
            int maxY       = iniY + hCell + 6;
            if (iniY >= maxBY - 3) continue;
            if (maxY > maxBY) maxY = maxBY;
            for (int j = 0; j < nCols; ++j)
            {
                const int iniX = minB + j * wCell;
                int maxX       = iniX + wCell + 6;
                if (iniX >= maxBX - 6) continue;
                if (maxX > maxBX) maxX = maxBX;
                CellDev c{};
                c.level = (short)l, c.x0 = (short)iniX, c.y0 = (short)iniY, c.w = (short)(maxX - iniX), c.h = (short)(maxY - iniY);
                c.off_x = (short)(j * wCell), c.off_y = (short)(i * hCell);
                {
                    const int a = iniX & 15;  // the TMA box starts at the 16-byte aligned column left of the ROI
                    c.box_h = 0, c.ld_inv = 0, c.ld_rows = 0;  // box_h is filled in once the level's tallest cell is known
                    const int zw = std::max(c.w - 6, 1), c0 = 3 + a, c1 = 3 + a + zw;
                    const int nwz = ((c1 + 3) >> 2) - (c0 >> 2);
                    c.zn_word0 = (unsigned short)(c0 >> 2), c.zn_words = (unsigned short)nwz;
                    c.zn_inv = (unsigned short)(65536 / nwz + 1), c.zn_rows = (unsigned short)(32 / nwz);
                }
                SNK_REQUIRE(c.w <= FAST_DIM && c.h <= FAST_DIM, "internal: cell %dx%d exceeds the tile", c.w, c.h);
                h->cells.push_back(c);
                // NMS survivors are strict local maxima inside the zone: no two are 8-neighbours
                const int zw = std::max(c.w - 6, 0), zh = std::max(c.h - 6, 0);
                cell_cap = std::max(cell_cap, ((zw + 1) / 2) * ((zh + 1) / 2));
            }
        }
        lv.n_cells  = (int)h->cells.size() - lv.cell_base;
        {
            int box_h = 1;
            for (size_t ci = lv.cell_base; ci < h->cells.size(); ++ci) box_h = std::max(box_h, (int)h->cells[ci].h);
            for (size_t ci = lv.cell_base; ci < h->cells.size(); ++ci) h->cells[ci].box_h = (unsigned short)box_h;
            h->box_h[l] = box_h;
            h->fast_rows = std::max(h->fast_rows, box_h);
            for (size_t ci = lv.cell_base; ci < h->cells.size(); ++ci)
                h->fast_list = std::max(h->fast_list, (std::max(h->cells[ci].w - 6, 0) * std::max(h->cells[ci].h - 6, 0) + 7) & ~7);
        }
        lv.sel_base = sel_acc;
        sel_acc += lv.quota + 4;
        for (int ty = 0; ty < (lv.h + BL_ROWS - 1) / BL_ROWS; ++ty)
            for (int tx = 0; tx < (lv.w + BL_TW - 1) / BL_TW; ++tx) h->tiles.push_back(TileDev{(short)l, (short)tx, (short)ty, 0});
        for (int band = 0; band < (lv.h + 4 + BM_BAND - 1) / BM_BAND; ++band)  // blocks start at row -4
            for (int strip = 0; strip < (lv.w + BM_TILE_W - 1) / BM_TILE_W; ++strip)  // strip = a warp's tile of 64 columns
                h->blur_tasks.push_back(BlurTask{(short)l, (short)strip, (short)band, 0});
    }
    P.cells_per_img = (int)h->cells.size();
    cell_cap        = align_up(cell_cap, 4);  // every cell list starts 16-byte aligned (k_distribute gathers with uint4 loads)
    P.cell_cap      = cell_cap;
    SNK_REQUIRE((size_t)P.cells_per_img * cell_cap < (1u << 30), "image too large for the candidate arrays");
    P.cand_per_img = P.cells_per_img * cell_cap;
    for (int l = 0; l < L; ++l) P.lv[l].cand_base = P.lv[l].cell_base * cell_cap;
    P.sel_per_img = sel_acc;
    P.kp_stride   = h->max_kp;

    int max_quota = 0;
    for (int l = 0; l < L; ++l) max_quota = std::max(max_quota, h->features[l]);
    h->node_cap = align_up(max_quota + 8, 4);  // multiple of 4 keeps the int2 array of the kernel 8-byte aligned
    SNK_REQUIRE(h->node_cap < (1 << gnusort::SHIFT), "nfeatures too large: %d nodes per level", h->node_cap);
    h->qt_smem = (size_t)h->node_cap * QT_NODE_BYTES + 2 * (size_t)QT_KP_SMEM * 4;
    {
        // the largest levels (those a noisy image can push beyond QT_KP_SMEM candidates) get a second launch with a larger
        // shared-memory capacity that only picks up the dense problems; SNK_QT_BIG_CAP / SNK_QT_BIG_LEVELS override for A/B runs
        const char* e1 = getenv("SNK_QT_BIG_CAP");
        const char* e2 = getenv("SNK_QT_BIG_LEVELS");
        h->qt_big_cap  = e1 ? atoi(e1) : QT_KP_SMEM_BIG;
        int nbig = 0;
        for (int l = 0; l < L; ++l)
            if ((long long)P.lv[l].w * P.lv[l].h >= 200000) nbig = l + 1;  // 752x480: levels 0 and 1
        h->qt_big_levels = e2 ? std::min(atoi(e2), L) : nbig;
        if (h->qt_big_cap <= QT_KP_SMEM) h->qt_big_levels = 0;
        h->qt_smem_big = (size_t)h->node_cap * QT_NODE_BYTES + 2 * (size_t)h->qt_big_cap * 4;
        if (h->qt_smem_big > (size_t)SMEM_CEILING) h->qt_big_levels = 0;
    }
    SNK_REQUIRE(20 * h->node_cap >= gnusort::cta_sort_scratch_bytes(h->node_cap), "internal: sort scratch does not fit");
    SNK_REQUIRE(h->qt_smem <= (size_t)SMEM_CEILING, "nfeatures too large for the distribution kernel's shared memory");
    // the attribute belongs to the kernel, not to this handle: every handle sets the same ceiling, so handles with
    // different image sizes (or threads creating handles at the same time) cannot lower it under each other
    SNK_CUDA(cudaFuncSetAttribute(k_distribute, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_CEILING));

    // + slack: the kernels read whole aligned words / short row segments that may extend a few bytes past a row's end
    SNK_CUDA(cudaMalloc(&h->d_pyr, pyr_bytes + 256));
    SNK_CUDA(cudaMalloc(&h->d_gauss, gauss_bytes + 256));
    for (int l = 0; l < L; ++l)
    {
        P.lv[l].img   = h->d_pyr + pyr_off[l];
        P.lv[l].gauss = h->d_gauss + gauss_off[l];
    }
    h->own_l0 = P.lv[0];
    for (int l = 0; l < L; ++l)
    {
        int rc = encode_level_map(h, l, P.lv[l].img, P.lv[l].pitch, P.lv[l].img_stride);
        if (rc) return rc;
    }
    h->map0_ptr = P.lv[0].img, h->map0_pitch = P.lv[0].pitch, h->map0_stride = P.lv[0].img_stride;
    if (h->blur_umma)
    {
        // per level: tensor maps, one task per column strip, the band matrices of its strips and row tiles
        SNK_CUDA(cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, h->prm.device));
        std::vector<UmmaBandJob> jobs;
        unsigned off = 0;
        for (int l = 0; l < L; ++l)
        {
            int rc = encode_umma_maps(h, l, P.lv[l].img, P.lv[l].pitch, P.lv[l].img_stride, true);
            if (rc) return rc;
            const int n_strips = (P.lv[l].w + UB_OUT_COLS - 1) / UB_OUT_COLS, n_row_tiles = (P.lv[l].h + UB_ROWS_OUT - 1) / UB_ROWS_OUT;
            const unsigned wr_off = off + (unsigned)n_strips * UB_WC_BYTES;
            for (int sidx = 0; sidx < n_strips; ++sidx)
            {
                h->umma_tasks.push_back(UmmaBlurTask{(short)l, (short)(sidx * UB_OUT_COLS), (short)P.lv[l].w, (short)P.lv[l].h, (short)n_row_tiles, (short)sidx,
                                                     (short)n_strips, (short)(sidx * UB_OUT_COLS >= 3 && sidx * UB_OUT_COLS + UB_OUT_COLS + 3 <= P.lv[l].w), off,
                                                     wr_off});
                h->umma_tiles_per_img += n_row_tiles;
                jobs.push_back(UmmaBandJob{off, (short)(sidx * UB_OUT_COLS), (short)P.lv[l].w, 0, 0});
                off += UB_WC_BYTES;
            }
            for (int t = 0; t < n_row_tiles; ++t)
            {
                jobs.push_back(UmmaBandJob{off, (short)(t * UB_ROWS_OUT), (short)P.lv[l].h, 1, 0});
                off += UB_WR_BYTES;
            }
        }
        UmmaBandJob* d_jobs = nullptr;
        SNK_CUDA(cudaMalloc(&h->d_umma_tasks, h->umma_tasks.size() * sizeof(UmmaBlurTask)));
        SNK_CUDA(cudaMemcpy(h->d_umma_tasks, h->umma_tasks.data(), h->umma_tasks.size() * sizeof(UmmaBlurTask), cudaMemcpyHostToDevice));
        SNK_CUDA(cudaMalloc(&h->d_umma_bands, off));
        SNK_CUDA(cudaMalloc(&d_jobs, jobs.size() * sizeof(UmmaBandJob)));
        SNK_CUDA(cudaMemcpy(d_jobs, jobs.data(), jobs.size() * sizeof(UmmaBandJob), cudaMemcpyHostToDevice));
        k_umma_bands<<<(unsigned)jobs.size(), 128>>>(d_jobs, h->d_umma_bands);
        SNK_CUDA(cudaGetLastError());
        SNK_CUDA(cudaDeviceSynchronize());
        SNK_CUDA(cudaFree(d_jobs));
        SNK_CUDA(cudaFuncSetAttribute(k_blur_umma, cudaFuncAttributeMaxDynamicSharedMemorySize, UB_SMEM));
    }
    {
        const char* e = getenv("SNK_RESIZE_TMA");
        h->resize_tma = !(e && e[0] == '0');
    }
    for (int l = 1; l < L; ++l)
    {
        std::vector<int> xo, yo;
        std::vector<short> xa, yb;
        resize_tables(P.lv[l - 1].w, P.lv[l].w, xo, xa, true);
        resize_tables(P.lv[l - 1].h, P.lv[l].h, yo, yb, false);
        // per-quad table of the resize kernel: 4 output pixels read a window of at most 8 source bytes
        const int sw = P.lv[l - 1].w, dw = P.lv[l].w;
        std::vector<ResizeQuad> quads((dw + 3) / 4);
        for (int qx = 0; qx < (int)quads.size(); ++qx)
        {
            ResizeQuad& q = quads[qx];
            const int x0  = 4 * qx;
            const int s0  = xo[x0];
            q.word        = s0 >> 2;
            q.shift_bits  = 8 * (s0 & 3);
            q._pad        = 0;
            unsigned nib[8];
            for (int j = 0; j < 4; ++j)
            {
                const int x  = std::min(x0 + j, dw - 1);  // lanes past the right edge repeat the last pixel (never stored)
                const int t0 = xo[x] - s0, t1 = std::min(xo[x] + 1, sw - 1) - s0;
                SNK_REQUIRE(t0 >= 0 && t1 <= 7, "scale factor %.3f not supported (max 1.3)", h->prm.scale_factor);
                nib[2 * j]     = (unsigned)t0;
                nib[2 * j + 1] = (unsigned)t1;
                q.coef[j]      = (unsigned)(unsigned short)xa[2 * x] | ((unsigned)(unsigned short)xa[2 * x + 1] << 16);
            }
            q.sel = 0;
            for (int k = 0; k < 8; ++k) q.sel |= nib[k] << (4 * k);
        }
        // per-row table: both source rows clamped as cv::resize's vertical pass clamps them, coefficients << 16
        const int sh = P.lv[l - 1].h, dh = P.lv[l].h;
        std::vector<ResizeRow> rows(dh);
        for (int y = 0; y < dh; ++y)
        {
            SNK_REQUIRE(yb[2 * y] >= 0 && yb[2 * y] <= 2048 && yb[2 * y + 1] >= 0 && yb[2 * y + 1] <= 2048,
                        "vertical coefficient outside [0, 2048]");
            rows[y].sy0 = std::min(std::max(yo[y], 0), sh - 1);
            rows[y].sy1 = std::min(std::max(yo[y] + 1, 0), sh - 1);
            rows[y].b0  = (unsigned)yb[2 * y] << 16;
            rows[y].b1  = (unsigned)yb[2 * y + 1] << 16;
        }
        // k_resize_tma: the largest source window of a 128 x 32 output tile (start rounded down to 16 bytes; a quad reads 3 words)
        {
            int bw = 16, bh = 1;
            for (int tx = 0; tx * (RT_W / 4) < (int)quads.size(); ++tx)
            {
                const int first = (quads[tx * (RT_W / 4)].word * 4) & ~15;
                const int lastq = std::min(tx * (RT_W / 4) + RT_W / 4 - 1, (int)quads.size() - 1);
                bw              = std::max(bw, quads[lastq].word * 4 + 12 - first);
            }
            for (int ty = 0; ty * RT_H < dh; ++ty)
                bh = std::max(bh, rows[std::min(ty * RT_H + RT_H - 1, dh - 1)].sy1 - rows[ty * RT_H].sy0 + 1);
            h->rs_box_w[l] = align_up(bw, 16), h->rs_box_h[l] = bh;
            if (h->rs_box_w[l] > 256 || bh > 256 || h->rs_box_w[l] * bh > 48 * 1024) h->resize_tma = false;  // beyond a TMA box: the plain kernel
        }
        SNK_CUDA(cudaMalloc(&h->d_quads[l], quads.size() * sizeof(ResizeQuad)));
        SNK_CUDA(cudaMalloc(&h->d_rows[l], rows.size() * sizeof(ResizeRow)));
        SNK_CUDA(cudaMemcpy(h->d_quads[l], quads.data(), quads.size() * sizeof(ResizeQuad), cudaMemcpyHostToDevice));
        SNK_CUDA(cudaMemcpy(h->d_rows[l], rows.data(), rows.size() * sizeof(ResizeRow), cudaMemcpyHostToDevice));
    }
    for (int l = 1; l < L && h->resize_tma; ++l)
    {
        int rc = encode_resize_map(h, l, P.lv[l - 1].img, P.lv[l - 1].pitch, P.lv[l - 1].img_stride);
        if (rc) return rc;
    }
    {
        const char* e   = getenv("SNK_DESCRIBE_TMA");
        h->describe_tma = !(e && e[0] == '0') && L <= 8;
        for (int l = 0; l < L && h->describe_tma; ++l)
        {
            int rc = encode_describe_maps(h, l, P.lv[l].img, P.lv[l].pitch, P.lv[l].img_stride, true);
            if (rc) return rc;
        }
        if (h->describe_tma) SNK_CUDA(cudaFuncSetAttribute(k_describe_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, DT_SMEM));
    }
    h->fast_rows = align_up(h->fast_rows, 8);  // keeps every tile buffer 128-byte aligned (8 rows x 96 or 176 bytes)
    h->fast_smem = 3 * (size_t)h->fast_rows * FAST_PITCH + 4 * (size_t)h->fast_list;
    SNK_REQUIRE(h->fast_smem <= (size_t)SMEM_CEILING, "internal: FAST tile too large");
    SNK_CUDA(cudaFuncSetAttribute(k_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_CEILING));
    // groups of horizontally adjacent cells (fast_group.cuh): same level and cell row, contiguous zones, at most
    // FASTG_CELLS cells and FASTG_GW ROI columns
    h->groups.clear();
    h->fastg_list = 8;
    {
        const char* e = getenv("SNK_FAST_CHUNK");
        h->fast_chunk = e && e[0] == '1';
    }
    for (size_t ci = 0; ci < h->cells.size();)
    {
        const CellDev& f = h->cells[ci];
        GroupDev g{};
        g.level = f.level, g.x0 = f.x0, g.y0 = f.y0, g.h = f.h, g.box_h = (short)f.box_h, g.off_x = f.off_x, g.off_y = f.off_y;
        g.cell0 = (int)ci;  // index in the image's cell list
        for (int k = 0; k <= FASTG_CELLS; ++k) g.zb[k] = 32767;
        int n = 0;
        while (ci + n < h->cells.size() && n < FASTG_CELLS)
        {
            const CellDev& c = h->cells[ci + n];
            if (c.level != f.level || c.y0 != f.y0 || c.h != f.h) break;
            const int rel = c.x0 - f.x0;
            if (n > 0 && rel + 3 != g.zb[n]) break;          // zones must touch
            if (rel + c.w > FASTG_GW) break;
            if (c.w < 7) break;                               // degenerate cell: keep it alone
            g.zb[n]     = (short)(rel + 3);
            g.zb[n + 1] = (short)(rel + c.w - 3);
            ++n;
        }
        if (n == 0)
        {  // a cell that cannot start a run (no zone): a group of its own with an empty zone
            n       = 1;
            g.zb[0] = 3, g.zb[1] = (short)std::max(3, f.w - 3);
        }
        g.ncell = (short)n;
        g.zend  = g.zb[n];
        const short zend = g.zend;
        for (int k = n + 1; k <= FASTG_CELLS; ++k) g.zb[k] = 32767;
        g.zb[n] = zend;
        {
            const int a = f.x0 & 15, c0 = a + g.zb[0], c1 = a + zend;
            if (h->fast_chunk)
            {  // 16-byte chunks of the tile rows that intersect the zone columns
                const int nq = std::max(((c1 + 15) >> 4) - (c0 >> 4), 1);
                g.zn_word0 = (unsigned short)(c0 >> 4), g.zn_words = (unsigned short)nq, g.zn_inv = (1u << 20) / nq + 1;
            }
            else
            {
                const int nwz = std::max(((c1 + 3) >> 2) - (c0 >> 2), 1);
                g.zn_word0 = (unsigned short)(c0 >> 2), g.zn_words = (unsigned short)nwz, g.zn_inv = (1u << 20) / nwz + 1;
            }
            SNK_REQUIRE(a + (zend + 3) <= FASTG_PITCH, "internal: cell group wider than the FAST tile");
        }
        h->fastg_list = std::max(h->fastg_list, (std::max(zend - g.zb[0], 0) * std::max(f.h - 6, 0) + 7) & ~7);
        h->groups.push_back(g);
        ci += n;
    }
    h->fastg_smem = 3 * (size_t)h->fast_rows * FASTG_PITCH + 4 * (size_t)h->fastg_list;
    SNK_REQUIRE(h->fastg_smem <= (size_t)SMEM_CEILING, "internal: FAST group tile too large");
    SNK_CUDA(cudaFuncSetAttribute(k_fast_group<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_CEILING));
    SNK_CUDA(cudaFuncSetAttribute(k_fast_group<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_CEILING));
    SNK_CUDA(cudaMalloc(&h->d_groups, h->groups.size() * sizeof(GroupDev)));
    SNK_CUDA(cudaMemcpy(h->d_groups, h->groups.data(), h->groups.size() * sizeof(GroupDev), cudaMemcpyHostToDevice));
    SNK_CUDA(cudaMalloc(&h->d_cells, h->cells.size() * sizeof(CellDev)));
    SNK_CUDA(cudaMemcpy(h->d_cells, h->cells.data(), h->cells.size() * sizeof(CellDev), cudaMemcpyHostToDevice));
    SNK_CUDA(cudaMalloc(&h->d_tiles, h->tiles.size() * sizeof(TileDev)));
    SNK_CUDA(cudaMemcpy(h->d_tiles, h->tiles.data(), h->tiles.size() * sizeof(TileDev), cudaMemcpyHostToDevice));
    SNK_CUDA(cudaMalloc(&h->d_blur_tasks, h->blur_tasks.size() * sizeof(BlurTask)));
    SNK_CUDA(cudaMemcpy(h->d_blur_tasks, h->blur_tasks.data(), h->blur_tasks.size() * sizeof(BlurTask), cudaMemcpyHostToDevice));
    const size_t cand_n = (size_t)B * P.cand_per_img;
    SNK_CUDA(cudaMalloc(&h->d_cand, cand_n * 4));
    SNK_CUDA(cudaMalloc(&h->d_workA, cand_n * 4));
    SNK_CUDA(cudaMalloc(&h->d_workB, cand_n * 4));
    SNK_CUDA(cudaMalloc(&h->d_cell_count, (size_t)B * P.cells_per_img * 4));
    SNK_CUDA(cudaMalloc(&h->d_cell_off, (size_t)B * P.cells_per_img * 4));
    SNK_CUDA(cudaMalloc(&h->d_sel, (size_t)B * P.sel_per_img * 4));
    SNK_CUDA(cudaMalloc(&h->d_sel_count, (size_t)B * L * 4));
    SNK_CUDA(cudaMalloc(&h->d_counts, (size_t)B * 4));
    SNK_CUDA(cudaMalloc(&h->d_err, 4));
    SNK_CUDA(cudaMemset(h->d_err, 0, 4));
    SNK_CUDA(cudaMalloc(&h->d_kps, (size_t)B * P.kp_stride * sizeof(snk_keypoint_f32)));
    SNK_CUDA(cudaMalloc(&h->d_desc, (size_t)B * P.kp_stride * 32));
    h->rows = rows;
    h->cols = cols;
    return SNK_OK;
}

extern "C" {

const char* snk_last_error(void) { return g_err; }
const char* snk_version(void) { return "snakeorb 0.1 (sm_100a)"; }

int snk_orb_create(const snk_orb_params* p, snk_orb** out)
{
    SNK_REQUIRE(p && out, "null argument");
    SNK_REQUIRE(p->nlevels >= 1 && p->nlevels <= SNK_MAX_LEVELS, "nlevels must be in 1..%d", SNK_MAX_LEVELS);
    SNK_REQUIRE(p->nfeatures >= 1, "nfeatures must be positive");
    SNK_REQUIRE(p->scale_factor > 1.0f && p->scale_factor <= 1.3f, "scale_factor must be in (1, 1.3]");
    SNK_REQUIRE(p->th_fast >= 1 && p->th_fast <= 254 && p->th_fast_min >= 0 && p->th_fast_min <= 254,
                "FAST thresholds out of range");
    SNK_REQUIRE(p->max_batch >= 1, "max_batch must be >= 1");
    int rc = select_device(p->device);
    if (rc) return rc;
    snk_orb* h = new snk_orb;
    h->prm     = *p;
    // ScalePyramid::ScalePyramid (saiga/vision/util/ScalePyramid.cpp:16-55); scale_factor is a float widened to double
    const int L       = p->nlevels;
    const double sf   = (double)p->scale_factor;
    const double fac  = 1.0 / sf;
    double per_level  = p->nfeatures * (1 - fac) / (1 - std::pow(fac, (double)L));
    int sum           = 0;
    h->scale[0] = h->inv_scale[0] = 1;
    for (int i = 0; i < L - 1; ++i)
    {
        h->features[i] = iround_f((float)per_level);
        sum += h->features[i];
        per_level *= fac;
    }
    for (int i = 1; i < L; ++i)
    {
        h->scale[i]     = sf * h->scale[i - 1];
        h->inv_scale[i] = 1 / h->scale[i];
    }
    h->features[L - 1] = std::max(p->nfeatures - sum, 0);
    h->max_kp          = 0;
    for (int i = 0; i < L; ++i) h->max_kp += h->features[i] + 3;  // Distribute returns at most quota + 3
    if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess)
    {
        set_error("cudaStreamCreate failed");
        delete h;
        return SNK_ERR_CUDA;
    }
    h->stream = h->own_stream;
    if (cudaStreamCreateWithFlags(&h->side_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_sync, cudaEventDisableTiming) != cudaSuccess ||
        cudaStreamCreateWithFlags(&h->dense_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_fork2, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_join2, cudaEventDisableTiming) != cudaSuccess ||
        cudaHostAlloc(&h->h_counts, sizeof(int) * p->max_batch, cudaHostAllocDefault) != cudaSuccess ||
        cudaHostAlloc(&h->h_err, sizeof(int), cudaHostAllocDefault) != cudaSuccess)
    {
        set_error("extractor setup failed (streams, events or %zu bytes of pinned memory): %s",
                  sizeof(int) * (size_t)p->max_batch, cudaGetErrorString(cudaGetLastError()));
        snk_orb_destroy(h);
        return SNK_ERR_CUDA;
    }
    *h->h_err = 0;
    // constants of the descriptor kernel
    int umax[16];
    {
        // ORBPattern::AngleUmax (saiga/vision/features/OrbPattern.cpp:16-36)
        const int R = 15;
        int vmax    = (int)std::floor(R * std::sqrt(2.f) / 2 + 1);
        int vmin    = (int)std::ceil(R * std::sqrt(2.f) / 2);
        for (int v = 0; v <= vmax; ++v) umax[v] = iround_f((float)std::sqrt((double)R * R - v * v));
        for (int v = R, v0 = 0; v >= vmin; --v)
        {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
    }
    static const int kUmax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
    if (memcmp(umax, kUmax, sizeof(umax)) != 0)
    {
        set_error("internal: AngleUmax table mismatch");
        snk_orb_destroy(h);
        return SNK_ERR_INVALID;
    }
    float2 pt[512];
    for (int i = 0; i < 512; ++i)  // entry j*32 + lane holds point 16*lane + j
        pt[(i & 15) * 32 + (i >> 4)] = make_float2((float)kOrbPattern[2 * i], (float)kOrbPattern[2 * i + 1]);
    if (cudaMalloc(&h->d_pattern_t, sizeof(pt)) != cudaSuccess ||
        cudaMemcpy(h->d_pattern_t, pt, sizeof(pt), cudaMemcpyHostToDevice) != cudaSuccess)
    {
        set_error("pattern upload failed: %s", cudaGetErrorString(cudaGetLastError()));
        snk_orb_destroy(h);
        return SNK_ERR_CUDA;
    }
    // word-wise moment weights: window byte p = 4 * w + b sits at column offset u = p - al - 15
    {
        std::vector<uint2> mt(DS_MOM_TABLE);
        for (int al = 0; al < 4; ++al)
            for (int r = 0; r < 31; ++r)
                for (int w = 0; w < DS_MOM_WORDS; ++w)
                {
                    const int v = r - 15, lim = umax[v < 0 ? -v : v];
                    uint2 e = make_uint2(0u, 0u);
                    for (int b = 0; b < 4; ++b)
                    {
                        const int u = 4 * w + b - al - 15;
                        if (u >= -lim && u <= lim)
                        {
                            e.x |= (unsigned)(unsigned char)(signed char)u << (8 * b);
                            e.y |= 1u << (8 * b);
                        }
                    }
                    mt[(al * 31 + r) * DS_MOM_WORDS + w] = e;
                }
        if (cudaMalloc(&h->d_moment_tab, mt.size() * sizeof(uint2)) != cudaSuccess ||
            cudaMemcpy(h->d_moment_tab, mt.data(), mt.size() * sizeof(uint2), cudaMemcpyHostToDevice) != cudaSuccess)
        {
            set_error("moment table upload failed: %s", cudaGetErrorString(cudaGetLastError()));
            snk_orb_destroy(h);
            return SNK_ERR_CUDA;
        }
    }
    *out = h;
    return SNK_OK;
}

void snk_orb_destroy(snk_orb* h)
{
    if (!h) return;
    cudaSetDevice(h->prm.device);
    cudaStreamSynchronize(h->stream);
    h->free_buffers();
    if (h->d_pattern_t) cudaFree(h->d_pattern_t);
    if (h->d_moment_tab) cudaFree(h->d_moment_tab);
    if (h->d_rgba) cudaFree(h->d_rgba);
    if (h->h_counts) cudaFreeHost(h->h_counts);
    if (h->h_err) cudaFreeHost(h->h_err);
    for (auto& set : h->ev)
        for (auto& e : set)
            if (e) cudaEventDestroy(e);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    if (h->side_stream) cudaStreamDestroy(h->side_stream);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->ev_sync) cudaEventDestroy(h->ev_sync);
    if (h->dense_stream) cudaStreamDestroy(h->dense_stream);
    if (h->ev_fork2) cudaEventDestroy(h->ev_fork2);
    if (h->ev_join2) cudaEventDestroy(h->ev_join2);
    delete h;
}

int snk_orb_set_stream(snk_orb* h, void* s)
{
    SNK_REQUIRE(h, "null handle");
    h->stream = s ? (cudaStream_t)s : h->own_stream;
    return SNK_OK;
}

int snk_orb_max_keypoints(const snk_orb* h) { return h ? h->max_kp : 0; }

int snk_orb_pyramid(const snk_orb* h, double* scale, int32_t* features)
{
    SNK_REQUIRE(h, "null handle");
    for (int i = 0; i < h->prm.nlevels; ++i)
    {
        if (scale) scale[i] = h->scale[i];
        if (features) features[i] = h->features[i];
    }
    return SNK_OK;
}

static void collect_set(snk_orb* h, int r)
{
    if (!h->ev_pending[r]) return;
    cudaEventSynchronize(h->ev[r][SNK_ORB_NUM_STAGES - 1]);
    float total = 0;
    for (int s = 0; s < SNK_ORB_NUM_STAGES - 1; ++s)
    {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, h->ev[r][s], h->ev[r][s + 1]) == cudaSuccess)
        {
            h->stage_ms[s] += ms;
            total += ms;
        }
        h->stage_launches[s] += h->ev_launches[r][s];
        h->stage_launches[SNK_ORB_NUM_STAGES - 1] += h->ev_launches[r][s];
    }
    h->stage_ms[SNK_ORB_NUM_STAGES - 1] += total;
    h->ev_pending[r] = false;
}
static void collect_events(snk_orb* h)
{
    for (int r = 0; r < snk_orb::PROF_RING; ++r) collect_set(h, r);
}

int snk_orb_enqueue(snk_orb* h, const uint8_t* imgs, int n_imgs, int rows, int cols, size_t pitch, size_t img_stride,
                    int on_device)
{
    SNK_REQUIRE(h, "null handle");
    SNK_REQUIRE(n_imgs >= 0 && n_imgs <= h->prm.max_batch, "n_imgs %d exceeds max_batch %d", n_imgs, h->prm.max_batch);
    SNK_CUDA(cudaSetDevice(h->prm.device));
    h->last_n = 0;
    if (n_imgs == 0 || !imgs || rows <= 0 || cols <= 0) return SNK_OK;  // ORBExtractor.cpp:182 — empty image: no output
    SNK_REQUIRE(pitch >= (size_t)cols, "pitch smaller than a row");
    SNK_REQUIRE(n_imgs == 1 || img_stride >= pitch * rows, "img_stride smaller than an image");
    int rc = orb_allocate(h, rows, cols);
    if (rc) return rc;
    PyrDev& P      = h->P;
    cudaStream_t st = h->stream;
    const bool prof = h->profiling;
    const int ring  = h->ev_next;
    int scratch_launches[SNK_ORB_NUM_STAGES] = {};
    int* nl         = prof ? h->ev_launches[ring] : scratch_launches;
    cudaEvent_t* ev = h->ev[ring];
    if (prof)
    {
        collect_set(h, ring);  // only blocks when the ring has wrapped onto a set that is still in flight
        memset(nl, 0, sizeof(scratch_launches));
        h->ev_next = (ring + 1) % snk_orb::PROF_RING;
        cudaEventRecord(ev[0], st);
    }

    nvtxRangePushA("snk_orb_enqueue");
    struct NvtxPop
    {
        ~NvtxPop() { nvtxRangePop(); }
    } nvtx_pop_enqueue;
    struct NvtxStage  // one open stage range at a time, closed on every return path
    {
        bool open = false;
        void next(const char* name)
        {
            if (open) nvtxRangePop();
            nvtxRangePushA(name);
            open = true;
        }
        ~NvtxStage()
        {
            if (open) nvtxRangePop();
        }
    } nvtx_stage;
    nvtx_stage.next("snk:level0");
    // stage 0: level 0. Device input is used in place (the reference copies it, ORBExtractor.cpp:269; nothing reads the
    // copy's apron, so aliasing is equivalent). Host input is copied straight into the level-0 buffer.
    // the kernels read rows as aligned 32-bit words and the FAST tiles through TMA (16-byte base / pitch / stride)
    const bool aligned = ((uintptr_t)imgs % 16 == 0) && (pitch % 16 == 0) && (img_stride % 16 == 0 || n_imgs == 1);
    if (on_device && aligned)
    {
        P.lv[0].img        = const_cast<uint8_t*>(imgs);
        P.lv[0].pitch      = (int)pitch;
        P.lv[0].img_stride = img_stride;
    }
    else
    {
        P.lv[0].img        = h->own_l0.img;
        P.lv[0].pitch      = h->own_l0.pitch;
        P.lv[0].img_stride = h->own_l0.img_stride;
        const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
        if (pitch == (size_t)P.lv[0].pitch && img_stride == P.lv[0].img_stride)
            SNK_CUDA(cudaMemcpyAsync(P.lv[0].img, imgs, img_stride * n_imgs, kind, st));
        else if (img_stride == pitch * rows)
            SNK_CUDA(cudaMemcpy2DAsync(P.lv[0].img, P.lv[0].pitch, imgs, pitch, cols, (size_t)rows * n_imgs, kind, st));
        else
            for (int i = 0; i < n_imgs; ++i)
                SNK_CUDA(cudaMemcpy2DAsync(P.lv[0].img + i * P.lv[0].img_stride, P.lv[0].pitch, imgs + i * img_stride,
                                           pitch, cols, rows, kind, st));
    }
    if (h->map0_ptr != P.lv[0].img || h->map0_pitch != P.lv[0].pitch || h->map0_stride != P.lv[0].img_stride)
    {
        const size_t stride0 = (n_imgs == 1 && P.lv[0].img_stride % 16) ? (size_t)P.lv[0].pitch * rows : P.lv[0].img_stride;
        rc = encode_level_map(h, 0, P.lv[0].img, P.lv[0].pitch, stride0);
        if (rc) return rc;
        if (h->blur_umma && (rc = encode_umma_maps(h, 0, P.lv[0].img, P.lv[0].pitch, stride0, false))) return rc;
        if (h->resize_tma && P.nlevels > 1 && (rc = encode_resize_map(h, 1, P.lv[0].img, P.lv[0].pitch, stride0))) return rc;
        if (h->describe_tma && (rc = encode_describe_maps(h, 0, P.lv[0].img, P.lv[0].pitch, stride0, false))) return rc;
        h->map0_ptr = P.lv[0].img, h->map0_pitch = P.lv[0].pitch, h->map0_stride = P.lv[0].img_stride;
    }
    if (prof) cudaEventRecord(ev[1], st);

    nvtx_stage.next("snk:pyramid");
    // stage 1: pyramid chain
    for (int l = 1; l < P.nlevels; ++l)
    {
        const LevelDev& s = P.lv[l - 1];
        const LevelDev& d = P.lv[l];
        // strip height: 8 output rows per warp — many short, independent chains suit this latency-bound kernel best
        // (measured on 512 images: 32/16/8 rows by level size 0.573 ms, 16 everywhere 0.554, 8 everywhere 0.552, 4: 0.590)
        const int rps = RS_ROWS / 4;
        dim3 grid((d.w + RS_TW - 1) / RS_TW, ((d.h + rps - 1) / rps + 3) / 4, n_imgs);
        if (h->resize_tma)
            k_resize_tma<<<dim3((d.w + RT_W - 1) / RT_W, (d.h + RT_H - 1) / RT_H, n_imgs), 128, (size_t)h->rs_box_w[l] * h->rs_box_h[l], st>>>(
                h->rs_map[l], d.img, d.img_stride, d.pitch, d.w, d.h, h->d_quads[l], h->d_rows[l], h->rs_box_w[l], h->rs_box_h[l], h->d_err);
        else if (l == 1 && s.img != h->own_l0.img)  // level 0 is the caller's device buffer: nothing may be read behind its last row
            k_resize<true><<<grid, 128, 0, st>>>(s.img, s.img_stride, s.pitch, s.h, d.img, d.img_stride, d.pitch, d.w, d.h,
                                                 h->d_quads[l], h->d_rows[l], rps);
        else
            k_resize<false><<<grid, 128, 0, st>>>(s.img, s.img_stride, s.pitch, s.h, d.img, d.img_stride, d.pitch, d.w, d.h,
                                                  h->d_quads[l], h->d_rows[l], rps);
        ++nl[1];
    }
    if (prof) cudaEventRecord(ev[2], st);

    nvtx_stage.next("snk:blur+fast");
    // blur of every level: on the side stream unless stages are being timed
    const bool overlap = !prof && h->side_stream && h->ev_fork && h->ev_join;
    if (overlap)
    {
        SNK_CUDA(cudaEventRecord(h->ev_fork, st));
        SNK_CUDA(cudaStreamWaitEvent(h->side_stream, h->ev_fork, 0));
        launch_blur(h, P, n_imgs, h->side_stream);
        SNK_CUDA(cudaEventRecord(h->ev_join, h->side_stream));
        ++nl[4];
    }

    // stage 2: FAST over every cell of every level
    if (h->per_cell_fast)
        k_fast<<<dim3((P.cells_per_img + FAST_CPB - 1) / FAST_CPB, n_imgs), FAST_NT, h->fast_smem, st>>>(
            P, h->maps, h->d_cells, h->d_cand, h->d_cell_count, h->d_err, h->fast_rows, h->fast_list);
    else
    {
        const dim3 grid(((int)h->groups.size() + FASTG_GPB - 1) / FASTG_GPB, n_imgs);
        if (h->fast_chunk)
            k_fast_group<true><<<grid, FASTG_NT, h->fastg_smem, st>>>(P, h->maps, h->d_groups, (int)h->groups.size(), h->d_cand,
                                                                      h->d_cell_count, h->d_err, h->fast_rows, h->fastg_list);
        else
            k_fast_group<false><<<grid, FASTG_NT, h->fastg_smem, st>>>(P, h->maps, h->d_groups, (int)h->groups.size(), h->d_cand,
                                                                       h->d_cell_count, h->d_err, h->fast_rows, h->fastg_list);
    }
    ++nl[2];
    if (prof) cudaEventRecord(ev[3], st);

    nvtx_stage.next("snk:distribute");
    // stage 3: quadtree distribution per (level, image)
    // the dense (level, image) problems run in their own launch (larger shared-memory array, fewer blocks per SM) on a
    // second stream beside the ordinary ones; on ordinary images that launch's blocks return at once
    if (h->qt_big_levels > 0)
    {
        cudaStream_t ds = prof ? st : h->dense_stream;
        if (!prof)
        {
            SNK_CUDA(cudaEventRecord(h->ev_fork2, st));
            SNK_CUDA(cudaStreamWaitEvent(ds, h->ev_fork2, 0));
        }
        k_distribute<<<dim3(n_imgs, h->qt_big_levels), QT_NT, h->qt_smem_big, ds>>>(
            P, h->d_cand, h->d_cell_count, h->d_cell_off, h->d_workA, h->d_workB, h->d_sel, h->d_sel_count, h->node_cap, h->qt_big_cap,
            h->d_err, 2, QT_KP_SMEM, h->qt_big_levels);
        if (!prof) SNK_CUDA(cudaEventRecord(h->ev_join2, ds));
        ++nl[3];
    }
    k_distribute<<<dim3(n_imgs, P.nlevels), QT_NT, h->qt_smem, st>>>(P, h->d_cand, h->d_cell_count, h->d_cell_off, h->d_workA,
                                                                     h->d_workB, h->d_sel, h->d_sel_count, h->node_cap, QT_KP_SMEM,
                                                                     h->d_err, h->qt_big_levels > 0 ? 1 : 0, QT_KP_SMEM, h->qt_big_levels);
    ++nl[3];
    if (h->qt_big_levels > 0 && !prof) SNK_CUDA(cudaStreamWaitEvent(st, h->ev_join2, 0));
    if (prof) cudaEventRecord(ev[4], st);

    // stage 4: blur (serial variant) / join
    if (overlap)
        SNK_CUDA(cudaStreamWaitEvent(st, h->ev_join, 0));
    else
    {
        launch_blur(h, P, n_imgs, st);
        ++nl[4];
    }
    if (prof) cudaEventRecord(ev[5], st);

    nvtx_stage.next("snk:describe");
    // stage 5: orientation + descriptors + output records
    if (h->describe_tma)
        k_describe_tma<<<dim3((P.kp_stride + DS_WARPS * DS_KP_PER_WARP - 1) / (DS_WARPS * DS_KP_PER_WARP), n_imgs), DS_WARPS * 32, DT_SMEM, st>>>(
            P, h->desc_maps, h->d_sel, h->d_sel_count, h->d_kps, h->d_desc, h->d_counts, h->d_pattern_t, h->d_moment_tab, h->d_err);
    else
        k_describe<<<dim3((P.kp_stride + DS_WARPS * DS_KP_PER_WARP - 1) / (DS_WARPS * DS_KP_PER_WARP), n_imgs), DS_WARPS * 32, 0, st>>>(
            P, h->d_sel, h->d_sel_count, h->d_kps, h->d_desc, h->d_counts, h->d_pattern_t, h->d_moment_tab);
    ++nl[5];
    if (prof)
    {
        cudaEventRecord(ev[6], st);
        cudaEventRecord(ev[7], st);
        h->ev_pending[ring] = true;
    }
    SNK_CUDA(cudaGetLastError());
    h->last_n = n_imgs;
    return SNK_OK;
}

int snk_orb_enqueue_rgba(snk_orb* h, const uint8_t* rgba, int n_imgs, int rows, int cols, size_t pitch, size_t img_stride,
                         int on_device)
{
    SNK_REQUIRE(h, "null handle");
    SNK_REQUIRE(n_imgs >= 0 && n_imgs <= h->prm.max_batch, "n_imgs %d exceeds max_batch %d", n_imgs, h->prm.max_batch);
    SNK_CUDA(cudaSetDevice(h->prm.device));
    h->last_n = 0;
    if (n_imgs == 0 || !rgba || rows <= 0 || cols <= 0) return SNK_OK;
    SNK_REQUIRE(pitch >= (size_t)cols * 4, "pitch smaller than a row of RGBA pixels");
    SNK_REQUIRE(n_imgs == 1 || img_stride >= pitch * rows, "img_stride smaller than an image");
    int rc = orb_allocate(h, rows, cols);
    if (rc) return rc;
    cudaStream_t st    = h->stream;
    const uint8_t* src = rgba;
    size_t spitch = pitch, sstride = img_stride;
    const bool aligned = ((uintptr_t)rgba % 16 == 0) && (pitch % 16 == 0) && (img_stride % 16 == 0 || n_imgs == 1);
    if (!on_device || !aligned)
    {
        // host input (or a device layout the 16-byte loads cannot read): dense copy into the staging buffer
        spitch  = ((size_t)cols * 4 + 15) & ~(size_t)15;
        sstride = spitch * rows;
        if (h->rgba_bytes < sstride * n_imgs)
        {
            SNK_CUDA(cudaStreamSynchronize(st));
            if (h->d_rgba) cudaFree(h->d_rgba);
            h->d_rgba = nullptr, h->rgba_bytes = 0;
            SNK_CUDA(cudaMalloc(&h->d_rgba, sstride * n_imgs));
            h->rgba_bytes = sstride * n_imgs;
        }
        const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
        if (img_stride == pitch * rows || n_imgs == 1)
            SNK_CUDA(cudaMemcpy2DAsync(h->d_rgba, spitch, rgba, pitch, (size_t)cols * 4, (size_t)rows * n_imgs, kind, st));
        else
            for (int i = 0; i < n_imgs; ++i)
                SNK_CUDA(cudaMemcpy2DAsync(h->d_rgba + i * sstride, spitch, rgba + i * img_stride, pitch, (size_t)cols * 4,
                                           rows, kind, st));
        src = h->d_rgba;
    }
    const LevelDev& l0 = h->own_l0;
    k_rgba_to_gray<<<dim3((cols + 511) / 512, rows, n_imgs), 128, 0, st>>>(src, sstride, spitch, l0.img, l0.img_stride,
                                                                          l0.pitch, cols, rows);
    SNK_CUDA(cudaGetLastError());
    // the internal level-0 buffer is 16-byte aligned: the extractor uses it in place
    return snk_orb_enqueue(h, l0.img, n_imgs, rows, cols, (size_t)l0.pitch, l0.img_stride, 1);
}

int snk_orb_device_results(snk_orb* h, const snk_keypoint_f32** kps, const uint64_t** desc, const int32_t** counts,
                           int* stride)
{
    SNK_REQUIRE(h && h->rows, "no results yet");
    if (kps) *kps = h->d_kps;
    if (desc) *desc = h->d_desc;
    if (counts) *counts = h->d_counts;
    if (stride) *stride = h->P.kp_stride;
    return SNK_OK;
}

int snk_orb_fetch(snk_orb* h, snk_keypoint_f32* kps, uint64_t* desc, int capacity, int* n_out)
{
    SNK_REQUIRE(h && n_out, "null argument");
    const int n = h->last_n;
    if (n == 0) return SNK_OK;
    SNK_REQUIRE(kps && desc && capacity >= 0, "null result arrays");
    SNK_CUDA(cudaSetDevice(h->prm.device));
    cudaStream_t st = h->stream;
    const int ks    = h->P.kp_stride;
    SNK_CUDA(cudaMemcpyAsync(h->h_counts, h->d_counts, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
    if (capacity >= ks && n > 1)
    {
        // whole-batch copies (rows of the device arrays are kp_stride records long)
        SNK_CUDA(cudaMemcpy2DAsync(kps, (size_t)capacity * sizeof(snk_keypoint_f32), h->d_kps,
                                   (size_t)ks * sizeof(snk_keypoint_f32), (size_t)ks * sizeof(snk_keypoint_f32), n,
                                   cudaMemcpyDeviceToHost, st));
        SNK_CUDA(cudaMemcpy2DAsync(desc, (size_t)capacity * 32, h->d_desc, (size_t)ks * 32, (size_t)ks * 32, n,
                                   cudaMemcpyDeviceToHost, st));
        SNK_CUDA(cudaStreamSynchronize(st));
    }
    else
    {
        SNK_CUDA(cudaStreamSynchronize(st));
        for (int i = 0; i < n; ++i)
        {
            const int c = h->h_counts[i];
            if (c > capacity)
            {
                set_error("image %d has %d keypoints, capacity is %d", i, c, capacity);
                return SNK_ERR_CAPACITY;
            }
            if (c == 0) continue;
            SNK_CUDA(cudaMemcpyAsync(kps + (size_t)i * capacity, h->d_kps + (size_t)i * ks, (size_t)c * sizeof(snk_keypoint_f32),
                                     cudaMemcpyDeviceToHost, st));
            SNK_CUDA(cudaMemcpyAsync(desc + (size_t)i * capacity * 4, h->d_desc + (size_t)i * ks * 4, (size_t)c * 32,
                                     cudaMemcpyDeviceToHost, st));
        }
        SNK_CUDA(cudaStreamSynchronize(st));
    }
    if (*h->h_err)
    {
        set_error("internal capacity error %d in the extraction kernels", *h->h_err);
        cudaMemsetAsync(h->d_err, 0, 4, st);
        return SNK_ERR_CAPACITY;
    }
    for (int i = 0; i < n; ++i) n_out[i] = h->h_counts[i];
    return SNK_OK;
}

int snk_orb_fetch_async(snk_orb* h, snk_keypoint_f32* kps, uint64_t* desc, int capacity, int* n_out)
{
    SNK_REQUIRE(h && kps && desc && n_out, "null argument");
    const int n = h->last_n;
    if (n == 0) return SNK_OK;
    const int ks = h->P.kp_stride;
    SNK_REQUIRE(capacity >= ks, "capacity %d is below snk_orb_max_keypoints() = %d", capacity, ks);
    SNK_CUDA(cudaSetDevice(h->prm.device));
    cudaStream_t st = h->stream;
    SNK_CUDA(cudaMemcpyAsync(n_out, h->d_counts, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    // the kernels' error flag is sticky on the device: the latest copy also shows earlier batches' faults (snk_orb_check)
    SNK_CUDA(cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
    if (capacity == ks)
    {
        SNK_CUDA(cudaMemcpyAsync(kps, h->d_kps, (size_t)n * ks * sizeof(snk_keypoint_f32), cudaMemcpyDeviceToHost, st));
        SNK_CUDA(cudaMemcpyAsync(desc, h->d_desc, (size_t)n * ks * 32, cudaMemcpyDeviceToHost, st));
    }
    else
    {
        SNK_CUDA(cudaMemcpy2DAsync(kps, (size_t)capacity * sizeof(snk_keypoint_f32), h->d_kps,
                                   (size_t)ks * sizeof(snk_keypoint_f32), (size_t)ks * sizeof(snk_keypoint_f32), n,
                                   cudaMemcpyDeviceToHost, st));
        SNK_CUDA(cudaMemcpy2DAsync(desc, (size_t)capacity * 32, h->d_desc, (size_t)ks * 32, (size_t)ks * 32, n,
                                   cudaMemcpyDeviceToHost, st));
    }
    return SNK_OK;
}

int snk_orb_check(snk_orb* h)
{
    SNK_REQUIRE(h, "null handle");
    if (!h->rows) return SNK_OK;  // nothing has run yet
    SNK_CUDA(cudaSetDevice(h->prm.device));
    SNK_CUDA(cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    SNK_CUDA(cudaStreamSynchronize(h->stream));
    if (*h->h_err)
    {
        set_error("internal capacity error %d in the extraction kernels", *h->h_err);
        *h->h_err = 0;
        SNK_CUDA(cudaMemsetAsync(h->d_err, 0, 4, h->stream));
        return SNK_ERR_CAPACITY;
    }
    return SNK_OK;
}

int snk_orb_last_batch(const snk_orb* h) { return h ? h->last_n : 0; }

int snk_orb_debug_set_error(snk_orb* h, int code)
{
    SNK_REQUIRE(h && h->rows, "no buffers yet: run one enqueue first");
    SNK_CUDA(cudaSetDevice(h->prm.device));
    SNK_CUDA(cudaMemcpyAsync(h->d_err, &code, sizeof(int), cudaMemcpyHostToDevice, h->stream));
    SNK_CUDA(cudaStreamSynchronize(h->stream));
    return SNK_OK;
}

int snk_orb_order_stream_after(snk_orb* h, void* stream)
{
    SNK_REQUIRE(h, "null handle");
    if ((cudaStream_t)stream == h->stream) return SNK_OK;
    SNK_CUDA(cudaSetDevice(h->prm.device));
    SNK_CUDA(cudaEventRecord(h->ev_sync, h->stream));
    SNK_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, h->ev_sync, 0));
    return SNK_OK;
}

int snk_orb_order_after_stream(snk_orb* h, void* stream)
{
    SNK_REQUIRE(h, "null handle");
    if ((cudaStream_t)stream == h->stream) return SNK_OK;
    SNK_CUDA(cudaSetDevice(h->prm.device));
    SNK_CUDA(cudaEventRecord(h->ev_sync, (cudaStream_t)stream));
    SNK_CUDA(cudaStreamWaitEvent(h->stream, h->ev_sync, 0));
    return SNK_OK;
}

int snk_orb_detect_batch(snk_orb* h, const uint8_t* imgs, int n_imgs, int rows, int cols, size_t pitch,
                         size_t img_stride, int on_device, snk_keypoint_f32* kps, uint64_t* desc, int capacity,
                         int* n_out)
{
    SNK_REQUIRE(h && n_out && kps && desc && capacity >= 0, "null argument");  // before any work is queued
    for (int i = 0; i < n_imgs; ++i) n_out[i] = 0;
    int rc = snk_orb_enqueue(h, imgs, n_imgs, rows, cols, pitch, img_stride, on_device);
    if (rc) return rc;
    return snk_orb_fetch(h, kps, desc, capacity, n_out);
}

int snk_orb_detect(snk_orb* h, const uint8_t* img, int rows, int cols, size_t pitch, int on_device,
                   snk_keypoint_f32* kps, uint64_t* desc, int capacity, int* n_out)
{
    return snk_orb_detect_batch(h, img, 1, rows, cols, pitch, pitch * (size_t)(rows > 0 ? rows : 0), on_device, kps,
                                desc, capacity, n_out);
}

int snk_orb_get_level(snk_orb* h, int img, int level, int blurred, const uint8_t** ptr, int* rows, int* cols,
                      size_t* pitch)
{
    SNK_REQUIRE(h && h->rows, "no image processed yet");
    SNK_REQUIRE(level >= 0 && level < h->prm.nlevels && img >= 0 && img < h->prm.max_batch, "index out of range");
    const LevelDev& L = h->P.lv[level];
    if (ptr) *ptr = blurred ? L.gauss + (size_t)img * L.gauss_stride : L.img + (size_t)img * L.img_stride;
    if (rows) *rows = L.h;
    if (cols) *cols = L.w;
    if (pitch) *pitch = blurred ? L.gauss_pitch : L.pitch;
    return SNK_OK;
}

int snk_orb_read_level(snk_orb* h, int img, int level, int blurred, uint8_t* dst, size_t dst_pitch)
{
    const uint8_t* p;
    int r, c;
    size_t pitch;
    int rc = snk_orb_get_level(h, img, level, blurred, &p, &r, &c, &pitch);
    if (rc) return rc;
    SNK_CUDA(cudaStreamSynchronize(h->stream));
    SNK_CUDA(cudaMemcpy2D(dst, dst_pitch, p, pitch, c, r, cudaMemcpyDeviceToHost));
    return SNK_OK;
}

int snk_orb_read_candidates(snk_orb* h, int img, int level, uint32_t* packed, int capacity, int* n_out)
{
    SNK_REQUIRE(h && h->rows && n_out, "no image processed yet");
    SNK_REQUIRE(level >= 0 && level < h->prm.nlevels && img >= 0 && img < h->prm.max_batch, "index out of range");
    SNK_CUDA(cudaStreamSynchronize(h->stream));
    const PyrDev& P   = h->P;
    const LevelDev& L = P.lv[level];
    std::vector<int> cnt(L.n_cells);
    SNK_CUDA(cudaMemcpy(cnt.data(), h->d_cell_count + (size_t)img * P.cells_per_img + L.cell_base, L.n_cells * 4,
                        cudaMemcpyDeviceToHost));
    int n = 0;
    for (int c = 0; c < L.n_cells; ++c)
    {
        if (n + cnt[c] <= capacity && cnt[c] > 0)
            SNK_CUDA(cudaMemcpy(packed + n, h->d_cand + ((size_t)img * P.cells_per_img + L.cell_base + c) * P.cell_cap,
                                cnt[c] * 4, cudaMemcpyDeviceToHost));
        n += cnt[c];
    }
    *n_out = n;
    return n <= capacity ? SNK_OK : SNK_ERR_CAPACITY;
}

int snk_orb_set_profiling(snk_orb* h, int enabled)
{
    SNK_REQUIRE(h, "null handle");
    collect_events(h);
    if (enabled && !h->ev[0][0])
    {
        SNK_CUDA(cudaSetDevice(h->prm.device));
        for (auto& set : h->ev)
            for (auto& e : set) SNK_CUDA(cudaEventCreate(&e));
    }
    h->profiling = enabled != 0;
    return SNK_OK;
}

int snk_orb_stage_ms(snk_orb* h, float* ms, int* launches, int reset)
{
    SNK_REQUIRE(h, "null handle");
    collect_events(h);
    for (int s = 0; s < SNK_ORB_NUM_STAGES; ++s)
    {
        if (ms) ms[s] = h->stage_ms[s];
        if (launches) launches[s] = h->stage_launches[s];
        if (reset) h->stage_ms[s] = 0, h->stage_launches[s] = 0;
    }
    return SNK_OK;
}

}  // extern "C"
