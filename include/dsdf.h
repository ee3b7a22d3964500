/*
 * dsdf.h -- C ABI of the B200-native DeepSDF decoder hot path (libdsdf_b200.so).
 *
 * Drop-in boundary for ONE path of anonymous-submission01/medical-shape-disentanglement:
 * forward/backward of networks/deep_sdf_decoder.Decoder (8x512, latent 256, latent_in=[4]) over
 * (latent, xyz) SDF samples, the auto-decoder optimisation step around it, and the grid decode
 * inside deep_sdf/mesh.create_mesh.  Plain pointers and sizes only; no torch types.  All device
 * pointers are CUDA device memory on the handle's device; every launch goes to the `stream`
 * argument (a cudaStream_t passed as void*), no hidden synchronisation, no allocation of caller
 * tensors.  Every entry returns 0 on success or a negative DSDF_E* code; the message is
 * available from dsdf_last_error().  There is no CPU fallback anywhere in the library.
 *
 * Each entry cites the reference interface it replaces (file:line under /root/reference).
 */
#ifndef DSDF_H_
#define DSDF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DSDF_OK 0
#define DSDF_EINVAL (-1)      /* bad argument / unsupported configuration            */
#define DSDF_ECUDA (-2)       /* CUDA runtime error (message has the CUDA string)    */
#define DSDF_ENOMEM (-3)      /* workspace too small / allocation failure            */
#define DSDF_EUNSUPPORTED (-4)/* valid reference config outside the fused path       */

#define DSDF_NUM_LIN 9

typedef struct dsdf_handle_s* dsdf_handle;

/* Network configuration == the NetworkSpecs the reference passes to
 * Decoder(latent_size, **specs["NetworkSpecs"]) (train_deep_sdf.py:348,
 * networks/deep_sdf_decoder.py:10-71).  Supported: latent_size 256, 8 hidden layers of 512,
 * latent_in == 4, xyz_in_all false, use_tanh false; weight_norm_mask bit l set <=> lin l is
 * wrapped by nn.utils.weight_norm (norm_layers with weight_norm=true). */
typedef struct {
  int32_t latent_size;
  int32_t hidden;
  int32_t num_hidden;
  int32_t latent_in;
  uint32_t weight_norm_mask;
  int32_t device;            /* CUDA device ordinal */
} dsdf_config;

/* Parameter addressing: one flat fp32 device buffer holding every decoder parameter with the
 * reference's shapes (state_dict of networks/deep_sdf_decoder.Decoder):
 *   w_off[l]: linL.weight [out,in]  (or linL.weight_v when weight-normalised)
 *   g_off[l]: linL.weight_g [out,1] (or -1)
 *   b_off[l]: linL.bias [out]
 * Offsets are in elements.  Gradients / Adam state use the same addressing. */
typedef struct {
  int64_t w_off[DSDF_NUM_LIN];
  int64_t g_off[DSDF_NUM_LIN];
  int64_t b_off[DSDF_NUM_LIN];
  int64_t total;             /* elements in the flat buffer */
} dsdf_param_layout;

/* Which specs["NetworkArch"] (train_deep_sdf.py:225) this build of the library implements: the one
 * source tree is built once per decoder family -- "deep_sdf_decoder" (libdsdf_b200.so,
 * networks/deep_sdf_decoder.py) and "siren_decoder" (libdsdf_b200_siren.so, -DDSDF_ARCH_SIREN,
 * networks/siren_decoder.py: sin(30x) activations, lin3 512->256, lin4 = [x3 | latent], no output
 * tanh).  Every entry below has the same signature and meaning in both. */
const char* dsdf_arch(void);

/* arch.Decoder(latent_size, **NetworkSpecs)  (train_deep_sdf.py:348) */
int dsdf_create(const dsdf_config* cfg, dsdf_handle* out);
int dsdf_destroy(dsdf_handle h);
const char* dsdf_last_error(dsdf_handle h);
/* number of SMs the persistent kernels are launched on (for workspace sizing / reports) */
int dsdf_num_sms(dsdf_handle h);

/* Fold the fp32 master parameters into the bf16 tensor-core images (weight-norm g/||v|| folded,
 * lin4 columns remapped, xyz/bias hi-lo split).  Must be called after every parameter update and
 * before forward/train calls.  Replaces the per-call nn.utils.weight_norm pre-forward hook
 * (networks/deep_sdf_decoder.py:49-53). */
int dsdf_set_weights(dsdf_handle h, const float* params, const dsdf_param_layout* layout,
                     void* stream);

/* Decoder.forward(input[N, L+3]) -> [N,1], inference (networks/deep_sdf_decoder.py:74-109). */
int dsdf_forward_dense(dsdf_handle h, const float* input, int64_t n, float* out, void* stream);

/* decode_sdf(decoder, latent_vector, queries) (deep_sdf/utils.py:86-97) generalised to
 * scene-contiguous segments: row r uses latent table[seg_ids[r / rows_per_seg]] and
 * xyz = xyz[r*xyz_stride + 0..2].  n_seg == 1, seg_ids == NULL means "one latent for all rows"
 * (table points at that latent). */
int dsdf_forward_segments(dsdf_handle h, const float* table, const int64_t* seg_ids,
                          int64_t n_seg, int64_t rows_per_seg, const float* xyz,
                          int64_t xyz_stride, int64_t n, float* out, void* stream);

/* create_mesh grid + decode loop (deep_sdf/mesh.py:38-71): SDF at linear grid indices
 * [lo, hi) of the N^3 lattice, coordinates generated on the fly bit-exactly as the reference's
 * fp32 torch expressions (including the true-division shear of mesh.py:46-47).  Sharding by
 * voxel slab == calling with disjoint [lo,hi). */
int dsdf_decode_grid(dsdf_handle h, const float* latent, int32_t grid_n, int64_t lo, int64_t hi,
                     float* out, void* stream);
/* the coordinates the grid decode uses (for bookkeeping tests): xyz[hi-lo][3] */
int dsdf_grid_coords(dsdf_handle h, int32_t grid_n, int64_t lo, int64_t hi, float* xyz,
                     void* stream);

/* Workspace (device bytes) needed by dsdf_train_fwd_bwd for up to max_rows rows in up to
 * max_segments segments. */
size_t dsdf_train_workspace_bytes(dsdf_handle h, int64_t max_rows, int64_t max_segments);

/* One fused gather + forward + clamped-L1 loss + code regulariser + backward of the
 * auto-decoder step body (train_deep_sdf.py:598-616,731) over n_seg scene-contiguous segments
 * of rows_per_seg rows (rows_per_seg % 128 == 0):
 *   sdf_data [n_seg*rows_per_seg, 4] = (x,y,z,sdf)            (train_deep_sdf.py:555-567)
 *   table    [num_scenes, 256] latent codes, rows seg_ids[b] are read (already max_norm-renormed:
 *            call dsdf_renorm_rows first, train_deep_sdf.py:429,598)
 *   n_total  global row count the loss is normalised by (train_deep_sdf.py:558,606; under
 *            scene sharding = sum over ranks)
 *   reg_lambda_ramp = CodeRegularizationLambda * min(1, epoch/100) (0 disables; :609-613)
 *   eikonal_lambda = 0.002 when specs["UseEikonal"] (train_deep_sdf.py:286,724-729), 0 disables: adds
 *            lambda * mean_rows (1 - ||d clamp(pred) / d xyz||)^2 (mean over n_total) and its gradient, computed
 *            without double backward: dgrad chain -> per-row (eps, xyz') -> the forward chain recomputed on the
 *            perturbed input with the first pass's ReLU masks -> one wgrad (ReLU decoder family only)
 * Outputs: grads (flat, dsdf_param_layout addressing, OVERWRITTEN), lat_grad [num_scenes,256]
 * (dense, OVERWRITTEN: zero for scenes not in the batch, as nn.Embedding(sparse=False)),
 * pred [n] (tanh output, un-clamped), loss_out[0] = L1 term, loss_out[1] = regulariser term,
 * loss_out[2] = Eikonal term, loss_out[3] = 0, or 1 + position of the first seg_id outside [0, num_scenes) (nn.Embedding raises on such
 * an index; here the step runs with that id replaced by 0 and reports it -- loss_out has 4 floats).
 * grads_ready_event (nullable cudaEvent_t): recorded on `stream` as soon as `grads` is final, before the
 * latent-gradient kernels: a scene-sharded caller starts its all-reduce of `grads` on a side stream there
 * (replaces nn.DataParallel's reduce_add of replica gradients, train_deep_sdf.py:352). */
int dsdf_train_fwd_bwd(dsdf_handle h, const float* params, const float* table,
                       int64_t num_scenes, const int64_t* seg_ids, int64_t n_seg,
                       int64_t rows_per_seg, const float* sdf_data, int64_t n_total,
                       float clamp_dist, float reg_lambda_ramp, float eikonal_lambda, void* workspace,
                       size_t workspace_bytes, float* grads, float* lat_grad, float* pred,
                       float* loss_out, void* grads_ready_event, void* stream);

/* nn.Embedding(max_norm) lookup side effect (train_deep_sdf.py:429,598): rows seg_ids[b] of
 * table[num_scenes,256] with norm > max_norm are scaled in place by max_norm/(norm+1e-7); ids outside
 * [0, num_scenes) are skipped (dsdf_train_fwd_bwd reports them). */
int dsdf_renorm_rows(dsdf_handle h, float* table, int64_t num_scenes, const int64_t* seg_ids,
                     int64_t n_seg, float max_norm, void* stream);

/* torch.nn.utils.clip_grad_norm_(params, max_norm) + torch.optim.Adam step
 * (train_deep_sdf.py:780-784, 445-454) over one flat buffer.  max_norm <= 0: no clipping.
 * scratch: >= 8 KiB device scratch; with max_norm > 0, scratch[0] = grad norm, scratch[1] = clip coef on
 * return (the norm is not computed when nothing is clipped).
 * `step` is the 1-based Adam step count. */
int dsdf_clip_adam(dsdf_handle h, float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                   int64_t n, float max_norm, float lr, float beta1, float beta2, float eps,
                   int64_t step, float* scratch, void* stream);

/* Decoder.forward(input[N, L+3]) under autograd (networks/deep_sdf_decoder.py:74-109 called with
 * grad enabled, e.g. reconstruct.py:90-94,130): arbitrary per-row latents, differentiable w.r.t. the
 * parameters AND the input.  forward_train = dsdf_forward_dense that also keeps the activation
 * stash + ReLU bit-masks in `workspace` (dsdf_dense_workspace_bytes(max_rows), ~2 MiB per 128 rows);
 * backward consumes that workspace with dOut[N] and writes grads (flat, dsdf_param_layout addressing,
 * OVERWRITTEN) and, if non-NULL, dInput[N, 259].  Tensor-core chain shared with the fused step;
 * the latent products that only exist for dense input run as fp32 tile GEMMs (dense_kernels.cuh). */
size_t dsdf_dense_workspace_bytes(dsdf_handle h, int64_t max_rows);
int dsdf_dense_forward_train(dsdf_handle h, const float* input, int64_t n, float* out, void* workspace,
                             size_t workspace_bytes, void* stream);
int dsdf_dense_backward(dsdf_handle h, const float* params, const float* input, int64_t n,
                        const float* out, const float* dout, void* workspace, size_t workspace_bytes,
                        float* grads, float* dinput, void* stream);

/* Stage-2 style call (train_MLP_VAE_deep_sdf.py:3097-3133: z_hat[B,256] produced by another network,
 * repeated over each shape's P samples, concatenated with xyz and decoded; gradients flow back into z_hat
 * and, if trainable, the decoder): the scene-structured form of the autograd path.  Row r of segment b uses
 * latents[b]; forward_train keeps stash + masks in `workspace` (dsdf_train_workspace_bytes(h, n, n_seg)),
 * backward turns dOut[n] into grads (flat, OVERWRITTEN) and dlatents[n_seg,256] (OVERWRITTEN) with the
 * deterministic segmented reduction.  rows_per_seg %% 128 == 0. */
int dsdf_segments_forward_train(dsdf_handle h, const float* params, const float* latents, int64_t n_seg,
                                int64_t rows_per_seg, const float* xyz, int64_t xyz_stride, float* out,
                                void* workspace, size_t workspace_bytes, void* stream);
int dsdf_segments_backward(dsdf_handle h, const float* params, const float* latents, int64_t n_seg,
                           int64_t rows_per_seg, const float* xyz, int64_t xyz_stride, const float* out,
                           const float* dout, void* workspace, size_t workspace_bytes, float* grads,
                           float* dlatents, void* stream);

/* Training-mode dropout of the hidden layers (networks/deep_sdf_decoder.py:24,103-104:
 * F.dropout(x, p=dropout_prob, training=self.training) after the ReLU of every layer in `dropout`),
 * fused into the forward epilogues of dsdf_train_fwd_bwd: layer_mask bit l <=> l in NetworkSpecs.dropout,
 * prob = dropout_prob.  Keep/drop of element (row, layer, column) is Philox4x32-10 keyed by `seed`
 * and counted by (row_base + row, layer, column block, offset): deterministic, order-independent;
 * pass a new `offset` (the step count) every step.  layer_mask 0 or prob 0 switches dropout off
 * (the default; eval mode).  Takes effect for subsequent dsdf_train_fwd_bwd calls on this handle. */
int dsdf_set_dropout(dsdf_handle h, uint32_t layer_mask, float prob, uint64_t seed, uint64_t offset,
                     int64_t row_base);
/* Bookkeeping view of that stream: keep_out[n_rows][512] (1 = kept) for rows [row_lo, row_lo+n_rows)
 * after lin `layer` under the current dsdf_set_dropout settings. */
int dsdf_dropout_keep(dsdf_handle h, int32_t layer, int64_t row_lo, int64_t n_rows, uint8_t* keep_out,
                      void* stream);

/* reconstruct(decoder, num_iterations, latent_size, test_sdf, stat, clamp_dist, num_samples, lr,
 * l2reg, code_reg_lambda, code_reg_type, code_bound, ..., dist_mean, dist_std, dist_weight,
 * dist_type) (reconstruct.py:16-151): Adam on the latent code of a FROZEN decoder, batched over
 * n_shapes independent shapes (each shape is exactly the reference's loop; shapes never interact).
 * The whole num_iterations loop is enqueued on `stream` by one call: per iteration a sample-window
 * gather (deep_sdf/data.py:112-136), the fused forward + clamped-L1(mean) + dgrad chain, the
 * per-shape latent gradient (segmented column sums, no atomics), Adam and the optional norm bound.
 * No wgrad, no activation stash.  lr schedule: lr * 0.1^(e // max(1, num_iterations/2)) (:35-44). */
typedef struct {
  int64_t num_iterations;
  int64_t num_samples;      /* rows per iteration and shape; even (half pos, half neg) */
  float clamp_dist;
  float lr;
  int32_t reg_kind;         /* 0 none; 1 lambda*mean(z^2) (l2reg / "l2_sq"); 2 lambda*||z|| ("l2_norm") */
  float reg_lambda;
  float code_bound;         /* <= 0: none; else z *= min(1, bound/(||z||+1e-12)) after every step */
  int32_t dist_kind;        /* 0 none; 1 mean(diff^2) ("zscore_l2"); 2 mean(|diff|) ("l1") */
  float dist_weight;
  const float* dist_mean;   /* device [256] or NULL */
  const float* dist_std;    /* device [256] (clamped >= 1e-8 by the caller) or NULL */
} dsdf_reconstruct_cfg;

size_t dsdf_reconstruct_workspace_bytes(dsdf_handle h, int64_t n_shapes, int64_t num_samples);

/*   samples     [total_rows, 4] (x,y,z,sdf): every shape's `pos` rows and `neg` rows (data.py:83-110)
 *   shape_base  [n_shapes, 2] int64 device: first row of the shape's pos block, of its neg block
 *   windows     [num_iterations, n_shapes, 2] int64 device: window starts inside pos / neg
 *               (the reference draws them with random.randint per iteration, data.py:118-131;
 *               the host side draws the same sequence and uploads it)
 *   latents     [n_shapes, 256] in/out; exp_avg / exp_avg_sq [n_shapes, 256] Adam state (zeroed by
 *               the caller for a fresh run); loss_hist [num_iterations, n_shapes] device (nullable);
 *   first_step  1-based Adam step / iteration index of the first iteration run by this call
 *               (1 for a fresh run; lets a caller split the loop into several calls). */
int dsdf_reconstruct(dsdf_handle h, const float* params, const dsdf_reconstruct_cfg* cfg,
                     int64_t n_shapes, const float* samples, const int64_t* shape_base,
                     const int64_t* windows, int64_t first_step, int64_t n_steps, float* latents,
                     float* exp_avg, float* exp_avg_sq, void* workspace, size_t workspace_bytes,
                     float* loss_hist, void* stream);

/* Per-section timing of dsdf_train_fwd_bwd with CUDA events recorded on the caller's stream (the stream the
 * kernels are launched on).  After enable(1) every call records an event at each section boundary; read()
 * synchronises on them and returns, summed over the n_calls recorded calls, the milliseconds of ms_sum[8]:
 * [0] scene fold + forward chain, [1] dgrad chain, [2] Eikonal recompute pass, [3] Eikonal row-sum corrections,
 * [4] wgrad GEMMs, [5] split-K reduction, [6] decoder-gradient small kernels, [7] latent-gradient small kernels +
 * loss.  enable(0/1) also clears the recorded events. */
int dsdf_profile_enable(dsdf_handle h, int32_t enable);
int dsdf_profile_read(dsdf_handle h, float* ms_sum, int32_t* n_calls);

/* Diagnostics (no reference counterpart): with DSDF_DBG=1 in the environment at dsdf_create the
 * persistent kernels accumulate per-CTA cycle counters (16 u64 per SM: MMA-issuer total / waits,
 * producer waits, epilogue waits).  Copies up to max_words counters to host memory `out`,
 * clears them, returns the number of words (or a negative error).  Synchronises the device. */
int dsdf_debug_read(dsdf_handle h, uint64_t* out, int32_t max_words);

/* Descriptor self-test: one-tile GEMMs through the same UMMA descriptor helpers.
 * kind 0: K-major  A[128x64] (1 SUB), B[128x64] (1 SUB) -> D[128][128] = A B^T
 * kind 1: MN-major A (2 SUBs), B (4 SUBs), K = rows 0..31 -> D[128][256] = A^T B */
int dsdf_selftest_gemm(int32_t kind, const void* a_img, const void* b_img, float* d_out,
                       void* stream);

/* ---- iso-surface extraction on the device (replaces the skimage.measure.marching_cubes call of the reference's
 * deep_sdf/mesh.py:119-121 `convert_sdf_samples_to_ply`; level / spacing as there).  No handle: plain kernels.
 * sdf = [nx, ny, nz] fp32, x slowest (the reference's sdf_values.reshape(N, N, N)).  Two passes:
 *   dsdf_mc_classify -> eflags[n] (bits 0..2: the point's +x/+y/+z edge crosses `level`) and
 *                       counts[n] = crossing edges | triangles of the point's cube << 32;
 *   the caller takes ONE inclusive prefix sum `scan` over counts (totals = scan[n-1]: V = low 32 bits, T = high),
 *   dsdf_mc_emit     -> verts[V,3] (grid index x spacing, linear interpolation along the edge; add the voxel-grid
 *                       origin as mesh.py:127-130 does) and faces[T,3] (int32 indices into verts).
 * inside = value < level; triangle normals point towards increasing values.  Vertex order: by owning grid point
 * (linear index), then edge axis.  The triangulation table is derived in tools/gen_mc_tables.py (face-consistent,
 * crack-free; not Lewiner's table -- see mc_kernels.cuh). */
int dsdf_mc_classify(const float* sdf, int32_t nx, int32_t ny, int32_t nz, float level, uint8_t* eflags,
                     int64_t* counts, void* stream);
int dsdf_mc_emit(const float* sdf, int32_t nx, int32_t ny, int32_t nz, float level, float spacing_x,
                 float spacing_y, float spacing_z, const uint8_t* eflags, const int64_t* scan, float* verts,
                 int32_t* faces, void* stream);

/* ---- SDF-sample generation on the device: CUDA restatement of the reference's src/PreprocessMesh.cpp (its OpenGL path
 * needs a graphics stack; this needs none).  No handle: plain kernels on raw device pointers.
 *  dsdf_rasterize_views: PreprocessMesh.cpp:443-477 + src/ShaderProgram.cpp -- orthographic 400 x 400 rasterisation of the
 *     (already centred) mesh from n_views camera poses; views = [n_views][12] rows of [R | t] (camera = R p + t, the
 *     ModelViewLookAt matrices); zbuf = n_views * 160000 uint64 scratch.  out_vert / out_norm [n_views][400][400][4] in
 *     the x-major order ValidPointsFromIm walks (src/Utils.cpp:28-43): (x, y, z, triangle id + 1) and the flat normal
 *     turned towards the camera; valid[...] = 1 where the reference's w != 0.  Compact with `valid` to get
 *     point_verts / point_normals.
 *  dsdf_sample_points: SampleFromSurface (:22-87) + the sample cloud of SampleSDFNearSurface (:104-135): n_surf surface
 *     points (area CDF `cdf[n_faces]` = inclusive sums of src/Utils.cpp:70-91 TriangleArea), two Gaussian perturbations
 *     each (std dev stdv1, stdv2), then n_rand uniform points in the cube of side cube_dim; Philox stream keyed by seed
 *     (the reference seeds from std::random_device).  xyz [2 n_surf + n_rand][3]; surf_out [n_surf][3] or NULL.
 *  dsdf_knn_sdf: SampleSDFNearSurface :137-172 -- for each xyz row the 11 nearest of the n_pts visible points (exact),
 *     |n . ray| if closer than stdv else |ray| from the nearest, sign by unanimous vote; sdf_out = NaN where the vote is
 *     split (the reference drops the sample).  pts / normals are float4 rows; nn_out [n][11] or NULL.
 *  dsdf_gather_windows: deep_sdf/data.py:112-136 for a batch of HBM-resident scenes: out[s] = cat(pos[ws:ws+half],
 *     neg[wn:wn+half]) from samples (float4 rows), base[s] = {pos row 0, neg row 0}, win[s] = {ws, wn}. */
int dsdf_rasterize_views(const float* verts, int64_t n_verts, const int32_t* faces, int64_t n_faces,
                         const float* views, int32_t n_views, uint64_t* zbuf, float* out_vert, float* out_norm,
                         uint8_t* valid, void* stream);
/* Allocation-free variant of dsdf_rasterize_views.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and results as dsdf_rasterize_views, but the depth buffer is a file-scope
 * static __device__ array inside the library (DSDF_RAST_STATIC_MAX_VIEWS x 400 x 400 x 8 B = 10.24 MB, held by every
 * GPU that loads the library until exit) instead of the caller's `zbuf`.  The call makes no cudaMalloc / cudaFree /
 * allocating library call.  n_views > DSDF_RAST_STATIC_MAX_VIEWS returns DSDF_EINVAL with a message naming the limit
 * (split the views over several calls).  All calls on one device share the array: it is reset at the start of each
 * call, and calls whose kernels overlap on the GPU (different streams without an ordering between them) race. */
#define DSDF_RAST_STATIC_MAX_VIEWS 8
int dsdf_rasterize_views_static(const float* verts, int64_t n_verts, const int32_t* faces, int64_t n_faces,
                                const float* views, int32_t n_views, float* out_vert, float* out_norm,
                                uint8_t* valid, void* stream);
int dsdf_sample_points(const float* verts, const int32_t* faces, const float* cdf, int64_t n_faces, int64_t n_surf,
                       int64_t n_rand, float stdv1, float stdv2, float cube_dim, uint64_t seed, float* xyz,
                       float* surf_out, void* stream);
int dsdf_knn_sdf(const float* xyz, int64_t n, const float* pts, const float* normals, int64_t n_pts, float stdv,
                 float* sdf_out, int32_t* nn_out, void* stream);
/* dsdf_knn_sdf through a uniform grid over the visible points (replaces the reference's nanoflann kd-tree,
 * PreprocessMesh.cpp:137-146): same arguments, same results bit for bit (exact neighbours, ties to the lower index), about
 * a thousand candidates per near-surface query instead of every point.  workspace: dsdf_knn_grid_workspace_bytes(n_pts)
 * bytes of device memory (the cell table and the cell-sorted copy of the points), rebuilt by every call. */
size_t dsdf_knn_grid_workspace_bytes(int64_t n_pts);
int dsdf_knn_sdf_grid(const float* xyz, int64_t n, const float* pts, const float* normals, int64_t n_pts, float stdv,
                      void* workspace, size_t workspace_bytes, float* sdf_out, int32_t* nn_out, void* stream);
int dsdf_gather_windows(const float* samples, const int64_t* base, const int64_t* win, int64_t n_scenes, int64_t half,
                        float* out, void* stream);

/* Diagnostics: tensor-pipe issue-rate micro-benchmark (iters tcgen05.mma M=128 N=256 K=16 on n_cta
 * CTAs); cycles_host[n_cta] = cycles per CTA.  kind bit0: SWIZZLE_128B B operand, bit1: commit per
 * MMA, bit2: N=128.  Synchronises the device. */
int dsdf_selftest_mma_rate(int32_t kind, int32_t iters, int32_t n_cta, uint64_t* cycles_host);

/* Diagnostics builds only (-DDSDF_HANG_TRAP): barrier waits that spin for seconds write
 * (block, thread, barrier smem address, parity) records into this mapped host buffer and trap.
 * buf[0] = record count, records at buf[4 + 4*i].  DSDF_EUNSUPPORTED in the product build. */
int dsdf_debug_hang_buffer(uint32_t** host_ptr);

/* Diagnostics builds only (-DDSDF_PROGRESS): every warp leader of the dgrad chain kernel posts its position
 * ((tile iteration << 16) | (layer << 8) | stage) before each barrier wait into this mapped host buffer
 * (word blockIdx * 64 + warp * 4, first 16 blocks), readable while a kernel hangs (tools/gpu_progress.py).
 * DSDF_EUNSUPPORTED in the product build. */
int dsdf_debug_progress_buffer(uint32_t** host_ptr);

#ifdef __cplusplus
}
#endif
#endif /* DSDF_H_ */
