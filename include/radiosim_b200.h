/* radiosim_b200 -- C ABI of the B200-native image-synthesis hot path.
 *
 * The reference (radionets-project/radiosim) is pure Python and has no FFI layer;
 * this library is the NEW layer underneath its generator functions.  Each entry
 * point below names the reference code whose pixel work it replaces
 * (paths relative to the reference checkout, e.g. src/radiosim/jets/jet.py:145).
 *
 * Conventions
 *  - plain C: pointers and sizes only, no C++/torch types.
 *  - every function returns 0 on success, a negative value for a CUDA error and a
 *    positive value for an argument error; rsb_last_error() describes the last
 *    failure on the calling thread.
 *  - "dev" pointers are device pointers on the current device, "host" pointers are
 *    host memory (pinned memory from rsb_malloc_host gives full PCIe bandwidth).
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *    Device-pointer entry points only enqueue work; they do not synchronise.
 *  - there is no CPU fallback: without a CUDA device every compute entry fails.
 */
#ifndef RADIOSIM_B200_H
#define RADIOSIM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RSB_ABI_VERSION 1

/* ------------------------------------------------------------------ runtime */
int rsb_abi_version(void);
const char* rsb_last_error(void);
int rsb_device_count(int* count);
int rsb_set_device(int device);
int rsb_get_device(int* device); /* the calling thread's current device */
/* Test aid without a reference counterpart: leaves `pattern` in every SM's shared memory, so that a
 * test can show a kernel does not depend on what its predecessor left there. */
int rsb_debug_scrub_shared(unsigned pattern, void* stream);
int rsb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor,
                    size_t* total_mem);
int rsb_malloc(void** dev_ptr, size_t bytes);
int rsb_free(void* dev_ptr);
int rsb_malloc_host(void** host_ptr, size_t bytes); /* pinned */
int rsb_free_host(void* host_ptr);
int rsb_memset(void* dev_ptr, int value, size_t bytes, void* stream);
int rsb_memcpy_h2d(void* dev_dst, const void* host_src, size_t bytes, void* stream);
int rsb_memcpy_d2h(void* host_dst, const void* dev_src, size_t bytes, void* stream);
/* rows x width_bytes with separate pitches (device planes into a host array that interleaves
 * them with other planes: the "gauss" targets of create_jet, reference jets/jet.py:219-225) */
int rsb_memcpy2d_d2h(void* host_dst, size_t dst_pitch, const void* dev_src, size_t src_pitch,
                     size_t width_bytes, size_t rows, void* stream);
int rsb_stream_create(void** stream);
int rsb_stream_destroy(void* stream);
int rsb_stream_sync(void* stream);
int rsb_device_sync(void);
int rsb_event_create(void** event);
int rsb_event_destroy(void* event);
int rsb_event_record(void* event, void* stream);
int rsb_stream_wait_event(void* stream, void* event); /* later work of `stream` waits for `event` */
int rsb_event_elapsed_ms(void* start, void* stop, float* ms); /* syncs on stop */
/* kernels launched by this library since load (all devices, this process) */
int rsb_launch_count(uint64_t* count);

/* Micro-benchmark: sustained MUFU.EX2 rate of the current device in ex2/s (best of
 * 5 launches of `iters` dependent ex2 x 8 chains per thread, CUDA events).  It is the
 * roofline denominator of the exp-bound splat kernels; no reference counterpart. */
int rsb_bench_sfu(int iters, double* ex2_per_second);

/* --------------------------------------------------------------------- jets */
/* Layout of one jet's parameters: `slots` components, each 6 float64 in the
 * order the reference hands to twodgaussian (src/radiosim/jets/jet.py:255-258):
 *   [amp, x (column centre), y (row centre), sx, sy, rotation].
 * A slot with amp == 0 is skipped (jet.py:252-253).  params is [n][slots][6]. */
#define RSB_JET_PARAM_STRIDE 6
#define RSB_JET_MAX_SLOTS 64

/* flags for rsb_jet_render* */
#define RSB_JET_NORMALIZE 1u /* conf.jet.scaling == "normalize" (jet.py:177-181) */
#define RSB_JET_OUT_F64 2u   /* out_img is float64 (the reference dtype), else float32 */

/* Bytes of device scratch rsb_jet_render needs for n images of size x size: 0 when the raw image stays
 * in shared memory -- up to 256 pixels (one CTA, or a 2-CTA cluster) and 401 .. 512 pixels (8-CTA cluster),
 * with at most 32 slots; otherwise one scratch image per resident CTA. */
int rsb_jet_workspace_bytes(int n, int slots, int size, size_t* bytes);

/* Render n jets -- replaces, per image, component_from_list + plane sum
 * (src/radiosim/jets/jet.py:145-146, :234-261, src/radiosim/utils/gauss.py:50-97),
 * zoom_on_source (src/radiosim/utils/utils.py:90-138), zoom_out (:141-177, image
 * only -- the reference leaves the planes untouched) and the max-normalisation
 * (jet.py:177-181).
 *
 *  params_dev   [n][slots][6] float64
 *  zoom_pad_dev [n] int32: < 0  -> the edge test (jet.py:149-157, evaluated by the
 *               host packer because it gates an RNG draw) failed: no zoom;
 *               >= 0 -> zoom branch, value = int64(pad_value) of jet.py:169-170.
 *  out_img_dev  [n][size][size] float32 (or float64 with RSB_JET_OUT_F64)
 *  out_margin_dev [n] int32: idx of zoom_on_source (utils.py:119), 0 when not zoomed
 *  out_peak_dev [n] float64: jet_max (jet.py:178) when RSB_JET_NORMALIZE, else 1.
 *               Evaluated in float64 at the float32 arg-max pixel.
 *  workspace_dev / workspace_bytes: >= rsb_jet_workspace_bytes(); may be NULL/0
 *               when that is 0.
 */
int rsb_jet_render(const double* params_dev, const int32_t* zoom_pad_dev, int n,
                   int slots, int size, unsigned flags, void* out_img_dev,
                   int32_t* out_margin_dev, double* out_peak_dev, void* workspace_dev,
                   size_t workspace_bytes, void* stream);

/* Jets of different sizes in ONE launch -- the nested create_jet calls of create_survey's
 * class-0 sources (src/radiosim/survey/survey.py:54-60: every sub-jet has its own
 * jet_size in [N/10.24, N/5.12)).  Same per-image work and results as rsb_jet_render.
 *
 *  sizes_dev    [n] int32: image i is sizes[i] x sizes[i], 2 <= sizes[i] <= max_size <= 256
 *  out_offs_dev [n] int64: image i starts at out_img_dev + out_offs[i] elements (a multiple
 *               of 4); its rows are (sizes[i] + 3) & ~3 elements apart (the pad columns are
 *               written with unspecified values)
 *  params_dev, zoom_pad_dev, out_margin_dev, out_peak_dev, flags: as for rsb_jet_render
 *  (slots <= 32).  Sorting the batch by size keeps the CTAs of one wave equally busy. */
int rsb_jet_render_var(const double* params_dev, const int32_t* zoom_pad_dev,
                       const int32_t* sizes_dev, const long long* out_offs_dev, int n, int slots,
                       int max_size, unsigned flags, void* out_img_dev, int32_t* out_margin_dev,
                       double* out_peak_dev, void* stream);

/* Component planes for training_type == "gauss" (jet.py:185, :219-225): the
 * `slots` planes after zoom_on_source (NOT zoomed out -- utils.py:168-175 drops
 * that result), divided by peak.  margin/peak come from rsb_jet_render.
 *  out_planes_dev [n][slots][size][size] float32/float64 (RSB_JET_OUT_F64) */
int rsb_jet_planes(const double* params_dev, const int32_t* zoom_pad_dev,
                   const int32_t* margin_dev, const double* peak_dev, int n, int slots,
                   int size, unsigned flags, void* out_planes_dev, void* stream);

/* Host-buffer convenience (the call a ctypes user makes): copies params in,
 * renders in chunks on internal streams with copy/compute overlap, copies images,
 * margins and peaks back, and synchronises.  Buffers should be pinned
 * (rsb_malloc_host) for full PCIe bandwidth; pageable memory works but is slower.
 * Allocates and caches device staging buffers on first use. */
int rsb_jet_render_host(const double* params_host, const int32_t* zoom_pad_host, int n,
                        int slots, int size, unsigned flags, void* out_img_host,
                        int32_t* out_margin_host, double* out_peak_host);

/* ------------------------------------------------------------------- survey */
/* Replaces the pixel work of create_survey (src/radiosim/survey/survey.py:49-122): every
 * gaussian / point source (survey.py:79-106) is a row [amp, x, y, sx, sy, rot] as handed to
 * twodgaussian (src/radiosim/utils/gauss.py:50-97); sub-jets (survey.py:54-76) are rendered
 * beforehand with rsb_jet_render into one packed float32 buffer and pasted here with the
 * reference's clipping; then the per-image max-normalisation and background plane
 * (survey.py:109-122).
 *
 *  src_params_dev [total_sources][6] float64, grouped per image: class-1 (gaussian) rows,
 *                 then class-2 (point) rows
 *  src_offs_dev   [n][3] int32: first class-1 row, first class-2 row, end row of image i
 *  jet_rects_dev  [jet_offs[n]] records {int64 pix; int32 size, row0, col0, pitch}: offset of the
 *                 sub-jet's first pixel in jet_pix_dev, its edge, the image position of
 *                 its pixel (0,0): row0 = floor(x - size/2), col0 = floor(y - size/2)
 *                 (survey.py:61-64; may be negative), and the floats between its rows
 *                 (size for rsb_jet_render output, (size + 3) & ~3 for rsb_jet_render_var)
 *  jet_offs_dev   [n + 1] int32: sub-jets of image i are jet_offs[i] .. jet_offs[i+1]-1
 *  out_x_dev      [n][size][size], out_y_dev [n][4][size][size]: float32, or float64 with
 *                 RSB_JET_OUT_F64 (planes: jets, gaussians, points, background)
 *  workspace_dev  >= rsb_survey_workspace_bytes() bytes of device scratch (coefficient records
 *                 and three counters per image; no image-sized scratch: the class planes stay on
 *                 chip until the image maximum is known and are written once, normalised)
 * Calls on different streams of one device are chained (each waits for the previous call's
 * kernel): the kernel is persistent and sized for the whole GPU. */
int rsb_survey_workspace_bytes(int n, int size, int total_sources, size_t* bytes);
int rsb_survey_render(const double* src_params_dev, const int32_t* src_offs_dev,
                      int total_sources, const void* jet_rects_dev, const int32_t* jet_offs_dev,
                      const float* jet_pix_dev, int n, int size, unsigned flags, void* out_x_dev,
                      void* out_y_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------- ppdisk */
/* Replaces the tensor work of create_proto (src/radiosim/ppdisks/ppdisks.py:95-185) on the
 * reference's fixed 900 x 900 grid.  The host replays the RNG chain and packs
 *
 *  comps_dev  [n_comps] records of 48 bytes:
 *             {float ca, sa;           float32(cos/sin(deg2rad(alpha)))       (:225-233)
 *              float a2_in, b2_in;     float32(a^2), float32(b^2) of the inner ellipse of a
 *                                      ring, a2_in <= 0 for the core          (:269, :305)
 *              float a2_out, b2_out;   the (outer) ellipse
 *              float scale;            _get_scale                              (:477-483)
 *              int32 sigma;            smoothing width                         (:135, :380-394)
 *              int32 image; int32 pad[3];}
 *  images_dev [n_images] records of 32 bytes:
 *             {int32 comp0, ncomp;     components of the image, core first
 *              int32 lo, hi;           crop [lo, hi) of the 900 grid           (:154-161)
 *              int32 new_size;         resize target                           (:166-171)
 *              int32 pad[3];}
 *  htab_dev   [17][900] float32: htab[s][t] = g_s(t - 449.5) / sum_t g_s(t - 449.5), the 1-D
 *             factor of astropy's Gaussian2DKernel(s, s, x_size=900, y_size=900)  (:446-450)
 *  out_dev    [n_images][img_size][img_size] float32 (the reference's output dtype)
 *
 * rsb_ppdisk_masks exposes the bit-exact mask step (_create_ellipse / _substract,
 * :188-271) alone: masks_dev is [n_comps][900][900] uint8. */
int rsb_ppdisk_workspace_bytes(int n_images, int n_comps, size_t* bytes);
int rsb_ppdisk_masks(const void* comps_dev, int n_comps, unsigned char* masks_dev, void* stream);
int rsb_ppdisk_render(const void* comps_dev, int n_comps, const void* images_dev, int n_images,
                      const float* htab_dev, int img_size, float* out_dev, void* workspace_dev,
                      size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------- mojave */
/* Replaces the array work of gen_compact / gen_one_jet / gen_two_jet
 * (src/radiosim/mojave/mojave.py:264-662): skewed-Gaussian jets (gauss.py:100-150, as pairs
 * of host-evaluated 1-D profiles), skimage.transform.swirl (mojave.py:259), rotated Gaussians
 * (gauss.py:50-97), skimage.transform.rotate (mojave.py:346,463,626), np.isclose zeroing and,
 * for compact sources, the arg-max recentre and rescale (mojave.py:628-635).
 *
 *  recipes_dev  [n] records of 224 bytes:
 *               {double g[3][6];            up to three [amp,x,y,sx,sy,rot] (amp 0 = unused)
 *                double strength[2], radius[2];   swirl of jet 0 / 1, centre (size/2, size/2)
 *                double angle;               rotation in degrees
 *                double amp;                 compact: final amplitude
 *                int32 blob0[2], nblob[2];   blobs of jet 0 / 1 in blobs_dev
 *                int32 flip1;                jet 1 is added flipped in both axes (:337)
 *                int32 compact; int32 pad[2];}
 *  blobs_dev    records of 16 bytes {float amp; int32 pad; int64 prof;}: prof is the offset
 *               (in floats) into profiles_dev of normal[size] followed by skew[size]
 *  out_dev      [n][size][size] float32, or float64 with RSB_JET_OUT_F64 */
int rsb_mojave_workspace_bytes(int n, int size, size_t* bytes);
int rsb_mojave_render(const void* recipes_dev, int n, const void* blobs_dev,
                      const float* profiles_dev, int size, unsigned flags, void* out_dev,
                      void* workspace_dev, size_t workspace_bytes, void* stream);

/* The host half of create_mojave in C++ (reference src/radiosim/mojave/mojave.py:56-64: one child
 * generator per image, rng.spawn; and the parameter draws of gen_compact :519-625, gen_one_jet
 * :426-461, gen_two_jet :292-344, gen_shocks :182-199, add_swirl :250-254): numpy's SeedSequence
 * spawning, PCG64 and the Generator distributions replayed bit for bit, scipy's expon / skewnorm
 * draws by their construction.  Host only.
 *   entropy [n_entropy], spawn_key [n_spawn_key] : the parent SeedSequence's entropy and spawn key
 *     as 32-bit words (little-endian words of the integers); children first_child ..
 *     first_child + n - 1 are the generators the reference's rng.spawn() calls would hand out
 *   kinds [n] : 0 gen_compact, 1 gen_one_jet, 2 gen_two_jet
 *   recipes [n] : 224-byte records as rsb_mojave_render takes them (blob0 = index into blob_params)
 *   blob_params [blob_cap] : 64-byte records {double x, y, amp, width, length, a; int32 ref_r, ref_c,
 *     pad[2]}: the skewed_gauss arguments of every blob; a shock carries in `amp` the factor drawn
 *     at mojave.py:190 and in (ref_r, ref_c) the pixel of the jet so far it multiplies
 *   *n_blobs_out : blobs written (an argument error if they exceed blob_cap) */
int rsb_mojave_pack(const uint32_t* entropy, int n_entropy, const uint32_t* spawn_key, int n_spawn_key,
                    long long first_child, int n, const int32_t* kinds, int size, void* recipes,
                    void* blob_params, int blob_cap, int32_t* n_blobs_out);
/* Device: blob_params -> the blob records and max-normalised 1-D profiles rsb_mojave_render reads
 * (gauss.py:131-148; shock amplitudes resolved blob after blob, mojave.py:190).
 *   blobs_out_dev [n_blobs] 16-byte records, profiles_out_dev [n_blobs][2][size] float32 */
int rsb_mojave_blobs(const void* recipes_dev, int n, const void* blob_params_dev, int n_blobs, int size,
                     void* blobs_out_dev, float* profiles_out_dev, void* stream);
/* add_swirl's image operation (mojave.py:259: skimage.transform.swirl(jet, center=, strength=, radius=),
 * other arguments at scikit-image's defaults: rotation 0, order 1, mode 'reflect', clip) on an arbitrary
 * float64 image in device memory.  center = (x, y) = (column, row); [clip_lo, clip_hi] = the input's
 * minimum and maximum.  Not in place. */
int rsb_swirl(const double* image_dev, int rows, int cols, double center_x, double center_y, double strength,
              double radius, double clip_lo, double clip_hi, double* out_dev, void* stream);
/* For tests: PCG64 (state, inc) of spawned children as numpy reports them, and raw draws of one
 * Generator method from such a state (kind 0 random, 1 standard_normal, 2 standard_exponential,
 * 3 power(a), 4 integers(a, b), 5 uniform(a, b)); state4 = {state hi, state lo, inc hi, inc lo}. */
int rsb_pcg64_spawn_states(const uint32_t* entropy, int n_entropy, const uint32_t* spawn_key, int n_spawn_key,
                           long long first_child, int n, uint64_t* out);
int rsb_pcg64_draws(uint64_t* state4, int kind, double a, double b, int count, double* out);

/* ---------------------------------------------------------------- utilities */
/* idx of zoom_on_source (src/radiosim/utils/utils.py:111-119): smallest distance from any
 * pixel > threshold to the border of a size x size float64 image; `size` if there is none
 * (the reference then uses idx = 0).  margin_dev: one int32. */
int rsb_source_margin(const double* img_dev, int size, double threshold, int* margin_dev,
                      void* stream);
/* cv2.resize(np.pad(src, pad), (dst_size, dst_size), INTER_LINEAR) on float64
 * (zoom_on_source utils.py:124-135 with pad = 0 on a crop, zoom_out utils.py:161-166).
 * `planes` images: plane p starts at src_dev + p * src_plane_stride, rows `src_pitch` apart;
 * dst_dev is [planes][dst_size][dst_size]. */
int rsb_resize_linear(const double* src_dev, int src_size, int src_pitch, size_t src_plane_stride,
                      int pad, int planes, double* dst_dev, int dst_size, void* stream);

/* rsb_jet_render without any allocation, for hosts that cannot call cudaMalloc once
 * they are running: a plugin inside a process whose framework has already reserved the
 * device for its own memory pool, or a service with a device-memory budget fixed at
 * start-up.  (No counterpart in the reference -- it exists only because of this
 * library's scratch requirement.)
 *
 * Same arguments, results and error convention as rsb_jet_render, minus the workspace:
 * for images too large for shared memory the per-CTA raw-image slots come from a 32 MiB
 * array compiled into the library.  That array occupies 32 MiB on every GPU that loads
 * librsb200.so, until the process exits.
 *  - sizes the on-chip renderer takes (up to 256) need no scratch: they run exactly as under
 *    rsb_jet_render and leave the array untouched.
 *  - limit: size <= 512 (larger sizes return an argument error; use rsb_jet_render).
 *    At 256x256 the array holds 128 slots, at 512x512 32, so fewer CTAs are resident
 *    than with a caller workspace; the images are identical.
 *  - the first call on each device looks up the array's device address (a one-time
 *    set-up, no allocation); make that first call before starting any stream capture.
 *  - all calls on a device share the one array: two calls whose kernels can overlap on
 *    the GPU (different streams, no ordering between them) race on it.  Calls on one
 *    stream, or otherwise ordered, are safe.  Nothing in it needs clearing between calls.
 *  - calls no cudaMalloc/cudaFree and no allocating library routine. */
int rsb_jet_render_static(const double* params_dev, const int32_t* zoom_pad_dev, int n,
                          int slots, int size, unsigned flags, void* out_img_dev,
                          int32_t* out_margin_dev, double* out_peak_dev, void* stream);

/* ---- host-side parameter packer (no GPU involved) -------------------------------------
 * Replaces the per-image RNG chain of create_jet, reference src/radiosim/jets/jet.py:54-125
 * (component count, amplitudes, knots, widths, rotations, boosting, centre shift, counter
 * jet) and the data-dependent draw at :149-168 (zoom_out factor iff the border maximum is
 * below 0.01), draw for draw on numpy's LEGACY global stream.
 *   mt_key[624], mt_pos, has_gauss, cached_gauss : the tuple of np.random.get_state(), updated
 *     in place -- write it back with np.random.set_state() and the stream continues where the
 *     reference's would;
 *   params [n, 2*comp_hi-1, 6] f64 (amp, x, y, sx, sy, rot), extra [n, 2*comp_hi-1, 2] f64
 *     (z_rotation, beta), zoom_pad [n] i32 (-1 or int64(pad_value), jet.py:169), zoom_out [n] f64
 *     : host buffers, as rsb_jet_render_host takes them;
 *   border_counts [3] or NULL: += images whose border test was decided by the upper bound, by
 *     one exact pixel, by the full border evaluation.
 * scaling="mojave" (a scipy beta draw per image) is not covered: callers keep the Python
 * packer for it.  1 <= comp_lo <= comp_hi <= 32. */
int rsb_jet_pack(uint32_t* mt_key, int32_t* mt_pos, int32_t* has_gauss, double* cached_gauss,
                 int n, int size, int comp_lo, int comp_hi, double* params, double* extra,
                 int32_t* zoom_pad, double* zoom_out, long long* border_counts);

/* The source loop of create_survey, reference src/radiosim/survey/survey.py:49-107 (class draw,
 * sub-jet size / position and its nested create_jet chain, wide and narrow Gaussians), on the same
 * legacy stream and with the same state hand-over as rsb_jet_pack.  class_dist [3] f64.  Buffers
 * sized for n*num_sources entries: src [.., 6] f64 (per image: wide rows, then narrow rows),
 * src_offs [n, 3] i32 (first wide, first narrow, end), and per sub-jet in draw order jet_params
 * [.., 2*comp_hi-1, 6] f64, jet_zoom_pad, jet_size, jet_image, jet_row0, jet_col0 (i32; corner of
 * the paste, survey.py:61-76); *njets_out = sub-jets drawn.  2 <= comp_lo <= comp_hi <= 32. */
int rsb_survey_pack(uint32_t* mt_key, int32_t* mt_pos, int32_t* has_gauss, double* cached_gauss,
                    int n, int size, int num_sources, const double* class_dist, int scale_sources,
                    int comp_lo, int comp_hi, double* src, int32_t* src_offs, double* jet_params,
                    int32_t* jet_zoom_pad, int32_t* jet_size, int32_t* jet_image, int32_t* jet_row0,
                    int32_t* jet_col0, int32_t* njets_out);

/* ---------------------------------------------------------------------- noise */
/* add_noise, reference src/radiosim/utils/utils.py:259-315 (called from
 * src/radiosim/simulations/simulations.py:101-108 for mojave bundles): a white field plus two
 * Gaussian-filtered fields (scipy fftconvolve with astropy's Gaussian2DKernel(kernel / 2),
 * kernels = [1, size/32, size/8]), each divided by its largest magnitude over the whole bundle and
 * weighted 0.2 / 0.3 / 0.5, the sum scaled per image to level * max(image) and added.
 *
 * The random fields are the reference's own draws, made by the host in the reference's order on
 * the legacy stream (rsb_legacy_normal_fill) and uploaded as float32:
 *  white_dev, field_a_dev, field_b_dev [n][size][size] float32: np.random.normal(size=shape) of
 *               utils.py:291 and of the two noise_small calls (:277)
 *  amp_a_dev, amp_b_dev [n] float32: max_noise = np.random.uniform(0, 1, n) of the two calls (:276)
 *  level_dev    [n] float64: np.random.uniform(*noise_level, size=n) / 100   (:306-308)
 *  image_dev, out_dev [n][size][size] float64 (out may be the image buffer itself)
 *  size <= 1024 (the column pass keeps a 32-column tile with its halo in shared memory). */
int rsb_noise_workspace_bytes(int n, int size, size_t* bytes);
int rsb_add_noise(const double* image_dev, const float* white_dev, const float* field_a_dev,
                  const float* field_b_dev, const float* amp_a_dev, const float* amp_b_dev,
                  const double* level_dev, int n, int size, double* out_dev, void* workspace_dev,
                  size_t workspace_bytes, void* stream);
/* count draws of np.random.normal() (legacy stream, loc 0, scale 1) as float32, in stream order:
 * the fields above (utils.py:277,291).  Host only; generator state as for rsb_jet_pack. */
int rsb_legacy_normal_fill(uint32_t* mt_key, int32_t* mt_pos, int32_t* has_gauss, double* cached_gauss,
                           long long count, float* out);

/* The RNG chain of generate_proto_set / create_proto (reference src/radiosim/ppdisks/ppdisks.py:74-76,
 * :135, :137, :305, :360-361, :474, :483) on numpy's legacy global stream, in C++ (host only; state
 * hand-over as rsb_jet_pack): n images -> component and image records for rsb_ppdisk_render.
 *   cos_tab, sin_tab [alpha_hi] : float32(cos / sin(deg2rad(alpha))) for every integer alpha, computed
 *     by the caller with numpy (the bit-exact masks are functions of these two numbers)
 *   comps [comp_cap] 48-byte records, images [n] 32-byte records; *n_comps_out = components written
 *     (an argument error if they exceed comp_cap; at most about 8 per image). */
int rsb_ppdisk_pack(uint32_t* mt_key, int32_t* mt_pos, int32_t* has_gauss, double* cached_gauss, int n,
                    int img_size, int alpha_lo, int alpha_hi, int ratio_lo, int ratio_hi,
                    double size_ratio_lo, double size_ratio_hi, const float* cos_tab,
                    const float* sin_tab, void* comps, int comp_cap, void* images, int32_t* n_comps_out);

#ifdef __cplusplus
}
#endif
#endif /* RADIOSIM_B200_H */
