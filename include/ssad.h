/*
 * ssad.h -- C ABI of libssad_gen.so: batched artificial-anomaly generation on B200 (sm_100a).
 *
 * The reference (gabry1998/Self-Supervised-Anomaly-Detection) has no FFI/plugin boundary: its
 * generator is plain Python (src/self_supervised/datasets.py:209-394,
 * src/self_supervised/dataset_generator.py:15-275) calling Pillow/torchvision per image.  This
 * header is the boundary a binding of that path would use: every entry point names the
 * reference function(s) it replaces.  INTEGRATION.md shows the ctypes binding.
 *
 * Conventions
 *  - all pointers except SsadBatchDesc* are DEVICE pointers owned by the caller (PyTorch);
 *    the library never allocates or frees device memory and never synchronises; work is ordered on `stream`
 *    (ssad_generate may fork part of a large batch onto internal helper streams and joins them back with events
 *    before it returns, so stream order and CUDA-graph capture see one call);
 *  - every random decision of the reference arrives as an explicit per-image parameter
 *    record (SsadImageParams); there is no RNG in the library;
 *  - images are uint8, interleaved RGB (HWC), row pitch in bytes;
 *  - masks are bit-packed: uint32 words, bit (x & 31) of word (x >> 5) of row y, 1 = object;
 *  - return value 0 = success, negative = SSAD_E* (ssad_strerror() gives the text).
 */
#ifndef SSAD_H_
#define SSAD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSAD_VERSION 100  /* 0.1.0 */

enum {
    SSAD_OK = 0,
    SSAD_EINVAL = -1,       /* bad argument (null pointer, non-positive size, bad enum) */
    SSAD_EALIGN = -2,       /* pointer / pitch not aligned as required */
    SSAD_EUNSUPPORTED = -3, /* valid request this build cannot serve (e.g. image too large for the fused path) */
    SSAD_ELAUNCH = -4,      /* CUDA launch / attribute error (cudaGetLastError text via ssad_last_cuda_error) */
    SSAD_EWORKSPACE = -5,   /* workspace too small */
    SSAD_ESTATIC = -6       /* request needs more scratch than the library's static device array holds
                               (ssad_obj_mask_static: SSAD_OBJ_MASK_STATIC_BYTES) */
};

/* mask_mode */
enum { SSAD_MASK_ALL = 0 /* textures: datasets.py:197-198 */, SSAD_MASK_SHARED = 1 /* fixed-pose classes: one
       cached mask, datasets.py:199-206 */, SSAD_MASK_PER_IMAGE = 2 /* hazelnut/screw/metal_nut, datasets.py:232-233 */ };

/* out_norm_dtype */
enum { SSAD_NORM_NONE = 0, SSAD_NORM_F32 = 1, SSAD_NORM_BF16 = 2 };

/* Per-image parameter record: every draw __getitem__ makes, with the
 * Python-float derivations already done on the host exactly as PIL / torchvision do them.
 * 192 bytes, 16-byte aligned array on the device. */
typedef struct SsadImageParams {
    int32_t  warp_fix[6];      /*   0: RandomAffine(3, scale=(1.05,1.1)) (datasets.py:221-222) as Pillow affine_fixed
                                       16.16 coefficients a0,a1,a2,a3,a4,a5 (half-pixel term folded into a2,a5) */
    int32_t  scar_fix[6];      /*  24: scar.rotate(angle, expand=True) (datasets.py:344), same encoding */
    uint32_t centre_r[6];      /*  48: get_random_coordinate (dataset_generator.py:259-265): index = (r * N) >> 32 */
    float    bright_f[2];      /*  72: the two ImageEnhance.Brightness factors (datasets.py:298-299), used iff cos > 0.99 */
    float    jit_f[4];         /*  80: ColorJitter factors brightness, contrast, saturation, (hue: unused) */
    int16_t  fallback_c[6][2]; /*  96: centre drawn inside the container when the mask is empty (dataset_generator.py:112-114) */
    int16_t  poly_xy[8][2];    /* 120: rect2poly vertices (dataset_generator.py:63-98), in draw order */
    int16_t  rect[4];          /* 152: generate_patch rectangle l,t,pw,ph inside the donor window (dataset_generator.py:172-191) */
    int16_t  scar_size[2];     /* 160: (nw, nh) of the rotated scar */
    int16_t  donor_off[2];     /* 164: x,y of the donor window (RandomCrop, datasets.py:253) inside the donor image */
    int16_t  crop_off[2];      /* 168: x,y of the working window inside src and mask (datasets.py:243-252) */
    int16_t  donor_idx;        /* 172: index into the donor bank (datasets.py:226); -1 = the image itself */
    uint8_t  label;            /* 174: 0 good, 1 polygon cut-paste, 2 scar, 3 long line (datasets.py:215,357-388) */
    uint8_t  warp_en;          /* 175 */
    uint8_t  src_variant;      /* 176: 0 cut, 1 average colour, 2 random colour (datasets.py:270-293) */
    uint8_t  scar_k;           /* 177: number of scar pastes, 2..5 (datasets.py:341) */
    uint8_t  poly_n;           /* 178: number of polygon vertices, 4..8 */
    uint8_t  rnd_rgb[3];       /* 179 */
    uint8_t  jit_order[4];     /* 182: ColorJitter fn_idx permutation (0 b, 1 c, 2 s, 3 hue/no-op) */
    uint8_t  jit_en;           /* 186: bit0 brightness, bit1 contrast, bit2 saturation; 0 = no jitter */
    uint8_t  line_n;           /* 187: label 3: points of the smoothed poly-line in SsadBatchDesc.line_xy (datasets.py:357-388) */
    uint32_t src_index;        /* 188: image of `src` this record works on when SsadBatchDesc.src_indexed is set */
} SsadImageParams;

/* Batch constants (host struct, passed by pointer, copied at call time). */
typedef struct SsadBatchDesc {
    int32_t batch;             /* images in this call */
    int32_t height, width;     /* working image x: imsize, or patch_size x patch_size in patch-level mode */
    int32_t src_height, src_width;
    int64_t src_pitch;         /* bytes per source row */
    int64_t src_stride;        /* bytes per source image */
    int32_t donor_count;       /* images in the donor bank (0: none) */
    int32_t donor_height, donor_width;
    int64_t donor_pitch, donor_stride;
    int32_t donor_win_h, donor_win_w; /* size of image_for_cutting as generate_patch sees it (== working size) */
    int32_t mask_mode;         /* SSAD_MASK_* */
    int32_t mask_pitch_words;  /* uint32 words per mask row */
    int64_t mask_stride_words; /* words per image (SSAD_MASK_PER_IMAGE) */
    int32_t mask_height, mask_width;
    int32_t box_poly[4];       /* Container(imsize, 1.75 | 1) left, top, right, bottom (dataset_generator.py:15-24) */
    int32_t box_scar[4];       /* Container(imsize, 2 | 1) */
    int32_t coverage_min3;     /* patch mode: label := 0 iff 3 * popcount(mask window) < coverage_min3 (datasets.py:258-259); 0 = off */
    int32_t max_patch_w, max_patch_h; /* upper bounds of rect[2], rect[3] over the batch (polygon table sizing) */
    int32_t out_norm_dtype;    /* SSAD_NORM_* */
    int32_t crop_align16;      /* 1: caller guarantees (3 * crop_off.x) % 16 == 0 for every image (lets TMA land the window dense) */
    int64_t out_u8_pitch;      /* bytes per output row (>= 3 * width) */
    int64_t out_u8_stride;     /* bytes per output image */
    int32_t any_warp;          /* 1: some records have warp_en set (sizes the workspace: x is materialised once per batch) */
    float   sat_factor_max;    /* upper bound of the enabled ColorJitter saturation factors jit_f[2] over the batch (the kernels'
                                  tables serve factors <= 2, torchvision's whole range for saturation <= 1) */
    int32_t src_indexed;       /* 1: record b reads source image params[b].src_index of a resident set of src_count images
                                  (the batch is never gathered: datasets.py:211 becomes an index); 0: image b of `src` */
    int32_t src_count;         /* images in `src` when src_indexed (SSAD_MASK_PER_IMAGE masks are then per SOURCE image:
                                  obj_mask depends on the original only, datasets.py:232-233) */
    int32_t line_width;        /* label 3: ImageDraw.line width, 3 image-level / 1 patch mode (datasets.py:383-388) */
    int32_t line_max;          /* points per image in line_xy */
    const int16_t* line_xy;    /* device [batch][line_max][2]: the poly-line vertices (x, y) of label-3 records, as ints the way
                                  ImageDraw.line truncates them; NULL when no record has label 3 */
    int32_t max_scar_w, max_scar_h; /* upper bounds of scar_size over the batch (sizes the rotated-scar colour canvas);
                                  0: use max_patch_w / max_patch_h.  With label-3 records max_patch_w / max_patch_h must
                                  also cover the bounding box of every poly-line (at most the working image). */
} SsadBatchDesc;

/* Optional per-image trace of the data-dependent decisions, for parity tests. */
typedef struct SsadDebug {
    int32_t n_coords;          /* popcount of the mask window (len(coords_map), datasets.py:264) */
    int32_t sim;               /* check_color_similarity(...) > 0.99 (datasets.py:295) */
    int32_t paste[5][2];       /* (left, top) of each paste (check_valid_coordinates_by_container) */
    int32_t contrast_mean;     /* ImageEnhance.Contrast degenerate grey level, -1 if contrast did not run */
    int32_t label_out;
    int32_t avg_rgb[3];        /* colour of the 'average' variant */
    double  cosine;
    int32_t reserved[2];
} SsadDebug;

int ssad_version(void);
const char* ssad_strerror(int code);
const char* ssad_last_cuda_error(void);
size_t ssad_sizeof_params(void);
size_t ssad_sizeof_desc(void);
size_t ssad_sizeof_debug(void);

/* The hot path: PretextTaskDataset.__getitem__ (datasets.py:209-394) for a whole batch, from the
 * resident uint8 originals to (x, y):  RandomAffine pre-warp -> patch-mode window -> mask-constrained
 * centre selection -> polygon cut-paste | rotated scar x k -> ColorJitter -> [ToTensor + Normalize].
 *   params     [batch] device records
 *   src        [batch] source images, uint8 HWC
 *   donor_bank [donor_count] donor images (textures; may be NULL when every donor_idx is -1)
 *   mask_bits  bit-packed mask(s) per mask_mode (NULL for SSAD_MASK_ALL)
 *   norm_lut   [3][256] table of Normalize(ToTensor(v)) per channel, element type per out_norm_dtype
 *              (NULL when out_norm is NULL)
 *   out_u8     [batch][height][width][3] uint8, nullable
 *   out_norm   [batch][3][height][width] f32/bf16 = Normalize(ToTensor(x)) (datasets.py:392-393), nullable
 *   y_out      [batch] int64 labels after the coverage rule, nullable
 *   debug      [batch] SsadDebug, nullable
 *   workspace  ssad_generate_workspace(desc) bytes, 256-byte aligned: per-image decisions and accumulators,
 *              polygon bit rows, and (any_warp) the warped images */
int ssad_generate(const SsadBatchDesc* desc, const SsadImageParams* params, const uint8_t* src,
                  const uint8_t* donor_bank, const uint32_t* mask_bits, const void* norm_lut,
                  uint8_t* out_u8, void* out_norm, int64_t* y_out, SsadDebug* debug,
                  void* workspace, size_t workspace_bytes, void* stream);

/* Contexts.  ssad_generate cuts a large batch into chunks that run side by side on helper streams; those streams and
 * their fork / join events belong to a context.  ssad_generate uses one default context per device (created on first
 * use under a lock), so calls from ONE host thread per device need nothing else.  Host threads that drive the same
 * device concurrently (the reference runs 8 loader workers, datasets.py:501-508) each create their own context for the
 * current device and call ssad_generate_ctx: the library keeps no other mutable state, so such calls are independent.
 * A context must be used on the device it was created on. */
typedef struct SsadContext SsadContext;
int ssad_context_create(SsadContext** out);
void ssad_context_destroy(SsadContext* ctx);
int ssad_generate_ctx(SsadContext* ctx, const SsadBatchDesc* desc, const SsadImageParams* params, const uint8_t* src,
                      const uint8_t* donor_bank, const uint32_t* mask_bits, const void* norm_lut,
                      uint8_t* out_u8, void* out_norm, int64_t* y_out, SsadDebug* debug,
                      void* workspace, size_t workspace_bytes, void* stream);
/* The library reads its tuning environment (SSAD_V7_*) once, at the first call; tests that change it call this. */
void ssad_reload_env(void);
size_t ssad_generate_workspace(const SsadBatchDesc* desc);
/* Number of kernels one ssad_generate call launches for this descriptor (the kernel pipeline cuts large batches into
 * chunks that run side by side on helper streams forked from, and joined back to, the caller's stream; a call that
 * is being captured into a CUDA graph -- captured != 0 -- uses more chunks than a plainly launched one). */
int ssad_generate_launches(const SsadBatchDesc* desc, int captured);

/* Host-computed constants of skimage.feature.canny(image, sigma=1.5, low_threshold=5, high_threshold=15) as
 * obj_mask calls it (dataset_generator.py:30).  They are inputs (not computed by the library) because the Gaussian
 * taps come from numpy's exp on the caller's machine -- scipy.ndimage._filters._gaussian_kernel1d(1.5, 0, 6) -- and
 * the masks must be bit-identical with the CPU libraries running there. */
typedef struct SsadCannyConsts {
    double gauss_w[13];        /* normalised taps exp(-x^2 / (2 sigma^2)), x = -6..6 */
    double low, high;          /* 5 / 255, 15 / 255 (uint8 input: thresholds are divided by the dtype max) */
    double eps;                /* numpy.finfo(float64).eps, added to the bleed-over divisor */
} SsadCannyConsts;

/* obj_mask (dataset_generator.py:27-39) for a batch: convert('L') -> canny(sigma 1.5, 5/255, 15/255) -> dilate 3x3
 * -> close 3x3 -> fill holes -> erode 4x4 -> largest 8-connected component, bit-packed (1 = object).
 * 13 <= width <= 1024, height >= 13.  workspace: ssad_obj_mask_workspace() bytes, 16-byte aligned. */
int ssad_obj_mask(int32_t batch, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                  int64_t src_stride, const SsadCannyConsts* consts /* host */, uint32_t* mask_bits,
                  int32_t mask_pitch_words, int64_t mask_stride_words, void* workspace, size_t workspace_bytes,
                  void* stream);
size_t ssad_obj_mask_workspace(int32_t batch, int32_t height, int32_t width);
/* Number of kernels one ssad_obj_mask call launches. */
int ssad_obj_mask_launches(void);
/* Test hook: ssad_obj_mask decides in float32 where a proven error margin allows and repeats the reference's float64
 * sequence elsewhere (per pixel, or per 64x32 tile when a tile has many such pixels).  The knobs widen the margins
 * (margin_scale > 1), cap the pixels a tile may list (sub_cap, 0..96) and the list entries per image (0..4096) so that
 * the tests can drive every path; the masks must not depend on them.  Arguments out of range restore the defaults. */
void ssad_obj_mask_test_knobs(float margin_scale, int32_t sub_cap, int32_t pix_per_image);
/* Test hook: the largest component is found by a union-find over the runs of set pixels when the plane holds at most
 * `max_runs` of them (default and maximum 4096), by peeling flooded components otherwise; 0 always takes the second
 * path, out of range restores the default.  The masks must not depend on it. */
void ssad_obj_mask_test_runs(int32_t max_runs);

/* ssad_obj_mask without a caller workspace and without any allocation.  (No counterpart in the reference: added
 * for callers that cannot allocate device memory -- code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc.)  Same arguments, same kernels and bit-identical masks as
 * ssad_obj_mask; the scratch (bit planes, undecided-pixel / tile lists, counters) lives in ONE file-scope
 * `static __device__` array of SSAD_OBJ_MASK_STATIC_BYTES inside the library, which every GPU that loads the library
 * holds until the process exits.  The call makes no cudaMalloc / cudaFree and calls no allocating library routine.
 *  - Limit: ssad_obj_mask_workspace(batch, height, width) <= SSAD_OBJ_MASK_STATIC_BYTES (16 MiB: e.g. 36 images of
 *    1024x1024, 290 of 256x256); a larger request returns SSAD_ESTATIC and launches nothing.
 *  - All calls on a device share the array.  It keeps the previous call's contents; the call clears the two counters
 *    it needs at zero itself, every other word is written before it is read.  Calls that overlap on the GPU (two
 *    streams, two host threads, or graphs replayed concurrently) race on it: serialise them, or use ssad_obj_mask
 *    with one workspace per stream. */
#define SSAD_OBJ_MASK_STATIC_BYTES (16u << 20)
int ssad_obj_mask_static(int32_t batch, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                         int64_t src_stride, const SsadCannyConsts* consts /* host */, uint32_t* mask_bits,
                         int32_t mask_pitch_words, int64_t mask_stride_words, void* stream);

/* Unit entry points (parity tests of the individual reference routines). */

/* rect2poly mask: ImageDraw.polygon(points, fill='white') on a pw x ph canvas (dataset_generator.py:99-100).
 * out: [n][max_h][max_w] uint8 0/255 (rows/cols beyond (pw, ph) are zero). */
int ssad_polygon_mask(int32_t n, const int16_t* poly_xy /* [n][8][2] */, const uint8_t* poly_n /* [n] */,
                      const int16_t* size /* [n][2] pw, ph */, uint8_t* out, int32_t max_w, int32_t max_h, void* stream);

/* ImageDraw.line(points, fill, width) without joints (datasets.py:383-388, label 3): n poly-lines of npts[i] <= 32 integer
 * vertices xy[i][k] = (x, y) (ImageDraw truncates the floats it is given) drawn with `line_width` (1: Bresenham segments
 * + the last point; > 1: one polygon per segment, Draw.c ImagingDrawWideLine) on width x height canvases.
 * out: [n][height][width] uint8 0 / 255.  All pointers device. */
int ssad_line_mask(int32_t n, const int16_t* xy, const uint8_t* npts, int32_t max_pts, int32_t line_width, uint8_t* out,
                   int32_t width, int32_t height, void* stream);

/* Image.transform(size, AFFINE, NEAREST) with affine_fixed coefficients (Image.rotate(expand=True) at
 * datasets.py:344, RandomAffine at datasets.py:221): channels in {3,4}; outside -> 0. */
int ssad_affine_nearest(const uint8_t* src, int32_t src_w, int32_t src_h, int32_t channels,
                        const int32_t* fix6_host, uint8_t* out, int32_t out_w, int32_t out_h, void* stream);

/* torchvision ColorJitter.forward (datasets.py:44-47,391) with explicit order/factors, in place semantics
 * on a batch of uint8 HWC images (dense pitch). order/factors are device arrays [n][4]. */
int ssad_color_jitter(int32_t n, int32_t height, int32_t width, const uint8_t* src, uint8_t* out,
                      const uint8_t* order, const float* factors, const uint8_t* enable, void* stream);

/* coords_map[k] (datasets.py:263-264 + dataset_generator.py:259-265): (x, y) of the k-th set pixel in row-major
 * order inside a window of a bit-packed mask; out_xy[i] = (-1,-1) when k >= count.  count_out nullable. */
int ssad_select_kth(const uint32_t* mask_bits, int32_t mask_pitch_words, int32_t win_x, int32_t win_y,
                    int32_t win_w, int32_t win_h, int32_t n, const uint32_t* k /* [n] device */,
                    int32_t* out_xy /* [n][2] device */, int32_t* count_out, void* stream);

/* paste_patch: image.paste(patch, (left, top), mask) (dataset_generator.py:268-275), in place on one dense uint8 RGB image.
 * patch [ph][pw][channels], channels in {3, 4}; mask [ph][pw] uint8 (0 keeps the image, 255 takes the patch, values in
 * between blend as Pillow's Paste.c does) or NULL:
 * then the alpha band of a 4-channel patch (the rotated scar pasted with itself as mask, datasets.py:354), else opaque.
 * The box is clipped to the image as Pillow does. */
int ssad_paste(uint8_t* image, int32_t width, int32_t height, const uint8_t* patch, int32_t pw, int32_t ph, int32_t channels,
               const uint8_t* mask, int32_t left, int32_t top, void* stream);

/* check_color_similarity (dataset_generator.py:147-159): cosine of the mean colours (float64, sklearn's normalise-then-dot
 * order) of two dense uint8 RGB images of n_image / n_patch pixels -> out[0] (device double).  The generator's own
 * decision (cosine > 0.99, datasets.py:295) is taken inside ssad_generate and reported in SsadDebug.cosine. */
int ssad_color_similarity(const uint8_t* image, int32_t n_image, const uint8_t* patch, int32_t n_patch, double* out, void* stream);

/* ToTensor(original) (datasets.py:394): uint8 HWC -> f32 NCHW / 255. */
int ssad_to_tensor(int32_t n, int32_t height, int32_t width, const uint8_t* src, int64_t src_pitch,
                   int64_t src_stride, float* out, void* stream);

/* Patch mode reads only a window of every source image (datasets.py:246-252) and, for labels 1 / 2 on non-texture classes,
 * the cut rectangle of image_for_cutting (datasets.py:253, dataset_generator.py:172-191).  When the images live in PINNED
 * host memory the device can address (cudaHostAlloc / torch pin_memory), this kernel fetches exactly those bytes over
 * PCIe into the device copy `dev_src` (same layout: pitch, stride), instead of a copy of the whole batch.  `params` are
 * the DEVICE records of the n images (image b of both buffers belongs to record b; crop_off, donor_off, rect, label,
 * donor_idx are read); max_patch_h bounds rect[3].  host_src, dev_src 16-byte aligned, src_stride a multiple of 16.
 * Not for records with warp_en (the pre-warp gathers from the whole image) or per-image obj_mask (it needs the whole
 * image on the device). */
int ssad_upload_windows(int32_t n, const SsadImageParams* params, const uint8_t* host_src, uint8_t* dev_src, int64_t src_pitch,
                        int64_t src_stride, int32_t src_height, int32_t src_width, int32_t win_height, int32_t win_width,
                        int32_t max_patch_h, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SSAD_H_ */
