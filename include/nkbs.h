/* nkbs.h — C ABI of the B200-native nerfacto / semantic-NeRF-W ray-batch path.
 *
 * Plain C: raw DEVICE pointers (unless marked HOST), sizes and scalars.  The caller (PyTorch) owns
 * every buffer; the library never allocates, frees or retains device memory, and never synchronises:
 * all work is enqueued on `stream` (a cudaStream_t passed as void*).  Every entry point returns 0 on
 * success and a negative code on error; nkbs_last_error() then returns a per-thread message.
 * Tensors are contiguous row-major fp32 unless stated; N = R*S samples, ray-major (n = r*S + s).
 *
 * Each entry replaces an operator of the third-party `nerfstudio` package (0.3.x, torch path) that
 * the reference reaches at the cited call sites (relative to the reference tree).  The reference has
 * no FFI of its own; INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 */
#ifndef NKBS_H
#define NKBS_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int nkbs_version(void);
const char* nkbs_last_error(void);

/* K1 — position front end + HashEncoding forward.
 * replaces: SceneContraction + HashEncoding.pytorch_fwd inside NerfactoField.get_density /
 *           HashMLPDensityField.get_density  (nerfacto.py:275-311,397-398; semantic_nerfw.py:102-126,187-188)
 * mode 0: x = normalised positions in [0,1]            (HashEncoding.forward itself)
 * mode 1: x = raw positions [N,3] -> contraction/aabb -> (x+2)/4 -> selector
 * mode 2: positions = origins[r] + directions[r]*(starts+ends)/2 from per-ray data, then as mode 1
 * bin_stride: floats between consecutive rays in starts / ends (0 = S; S+1 when both are views of one [R,S+1]
 * bin-edge array, starts = bins, ends = bins + 1).  The same convention holds for K6/K7/K8.
 * aabb, scalings: HOST pointers ([6] and [L]).  xn/selector/indices may be NULL.
 * indices: int64 [N,L,8] table rows in the torch path's corner order (parity checks). */
int nkbs_encode_fwd(int32_t mode, const float* x, const float* origins, const float* directions,
                    const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S, int32_t contraction,
                    const float* aabb, const float* table, const float* scalings, int32_t L, int32_t log2_T,
                    int32_t F, float* xn, float* selector, float* features, int64_t* indices, void* stream);

/* K2 — HashEncoding backward: scatters dfeat [N,2L] into dtable [L*T,2] (accumulates).
 * replaces: autograd of hash_table[hashed_k] gathers (same call sites as K1).
 * dedup_levels: levels [0,dedup_levels) merge duplicate rows inside a warp before the vector
 * reduction; <0 = all levels. */
int nkbs_encode_bwd(const float* xn, const float* dfeat, int64_t N, const float* scalings, int32_t L,
                    int32_t log2_T, int32_t F, int32_t dedup_levels, float* dtable, void* stream);

/* K3 (fp32) — fused MLP of 1..3 nn.Linear layers (torch [out,in] weights), widths <= 64.
 * replaces: field_components.mlp.MLP + field-head activations (nerfacto.py:275-311,398;
 *           semantic_nerfw.py:102-126,188).
 * dims: HOST [nl+1]; W,b,dW,db: HOST arrays of nl DEVICE pointers.  relu_mask bit l = ReLU after layer l.
 * out_act: 0 none, 1 sigmoid, 2 density (ch0 = exp(raw)*selector, trunc_exp backward), 3 transient
 * heads (ch0 softplus, ch1-3 sigmoid, ch4 softplus).  selector [N] may be NULL.
 * bwd recomputes the forward from x; dx may be NULL; dW/db are accumulated into. */
int nkbs_mlp_fwd(const float* x, int64_t N, int32_t nl, const int32_t* dims, const float* const* W,
                 const float* const* b, int32_t relu_mask, int32_t out_act, const float* selector, float* y,
                 float* y0, void* stream);
int nkbs_mlp_bwd(const float* x, const float* dy, int64_t N, int32_t nl, const int32_t* dims,
                 const float* const* W, const float* const* b, int32_t relu_mask, int32_t out_act,
                 const float* selector, const float* dy0, float* dx, float* const* dW, float* const* db,
                 void* stream);
/* y0 (may be NULL): column 0 of y written once more as a contiguous [N] array — torch.split(h, [1, geo_feat_dim])
 * of NerfactoField.get_density without a copy kernel; dy0 (may be NULL): [N], added to column 0 of dy inside the
 * kernel (the gradient that arrives through y0) instead of a separate add kernel.  dW / db are ACCUMULATED into. */

/* K3 (tensor cores) — the same fused MLP on tcgen05 tensor cores: 16-bit operands (fmt 0 = fp16, 1 = bf16),
 * fp32 accumulation in TMEM, weights resident in shared memory, bias folded into the GEMM.  Same argument
 * meaning as nkbs_mlp_fwd/bwd.  This is the mixed-precision path (the reference trains under autocast,
 * train.py:50); parity bar 1e-2 relative. */
int nkbs_mlp_tc_fwd(const float* x, int64_t N, int32_t nl, const int32_t* dims, const float* const* W,
                    const float* const* b, int32_t relu_mask, int32_t out_act, int32_t fmt,
                    const float* selector, float* y, float* y0, void* stream);
int nkbs_mlp_tc_bwd(const float* x, const float* dy, int64_t N, int32_t nl, const int32_t* dims,
                    const float* const* W, const float* const* b, int32_t relu_mask, int32_t out_act,
                    int32_t fmt, const float* selector, const float* dy0, float* dx, float* const* dW,
                    float* const* db, void* stream);

/* K8 — the proposal density field fused: ray point -> contraction -> HashEncoding -> MLP(2L -> H -> 1, ReLU) ->
 * trunc_exp * selector in one kernel per direction; only the density [N] is written.  fp32 throughout.
 * replaces: HashMLPDensityField.get_density / density_fn (nerfacto.py:299-311,325-332; semantic_nerfw.py:119-137).
 * mode 1: x = raw positions [N,3] (R = N, S = 1); mode 2: per-ray origins/directions + per-sample starts/ends.
 * Supported shapes: (L=5,H=16) nerfacto proposals, (L=8,H=64) HashMLPDensityField defaults; others return an
 * error (callers use K1 + K3).  W1 [H,2L], b1 [H], W2 [1,H], b2 [1] in torch nn.Linear layout.
 * bwd recomputes the forward and ACCUMULATES into dtable [L*T,2], dW1, db1, dW2, db2. */
int nkbs_density_fwd(int32_t mode, const float* x, const float* origins, const float* directions,
                     const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S, int32_t contraction,
                     const float* aabb, const float* table, const float* scalings, int32_t L, int32_t log2_T,
                     const float* W1, const float* b1, const float* W2, const float* b2, int32_t H,
                     float* density, void* stream);
/* nkbs_density_fwd that also leaves what a backward pass through the UNFUSED kernels needs: feat_out [N, 2L] hash-grid
 * features (16-byte aligned), xn_out [N,3] normalised positions x selector, sel_out [N] selector.  The hidden-64 field
 * trains this way: one fused forward launch, then nkbs_mlp_tc_bwd / nkbs_mlp_bwd (x = feat_out, out_act = density,
 * selector = sel_out) and nkbs_encode_bwd (xn = xn_out) — its fused backward is register-bound.
 * xn_out may be NULL (the ray-major scatter recomputes the positions).
 * nkbs_density_bwd_saved: backward of the (L=5, H=16) field from those saved features and selectors — no positions, no
 * table gathers: reads feat_in [N,2L], sel_in [N], ddensity [N], accumulates dW1 / db1 / dW2 / db2 and writes the
 * feature gradients dfeat_out [N,2L] (zero where the selector is zero); nkbs_encode_bwd_rays / nkbs_encode_bwd
 * scatter them. */
int nkbs_density_fwd_save(int32_t mode, const float* x, const float* origins, const float* directions,
                          const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S, int32_t contraction,
                          const float* aabb, const float* table, const float* scalings, int32_t L, int32_t log2_T,
                          const float* W1, const float* b1, const float* W2, const float* b2, int32_t H,
                          float* density, float* feat_out, float* xn_out, float* sel_out, void* stream);
int nkbs_density_bwd_saved(int64_t N, int32_t L, int32_t H, const float* feat_in, const float* sel_in,
                           const float* ddensity, const float* W1, const float* b1, const float* W2, const float* b2,
                           float* dW1, float* db1, float* dW2, float* db2, float* dfeat_out, void* stream);
int nkbs_density_bwd(int32_t mode, const float* x, const float* origins, const float* directions,
                     const float* starts, const float* ends, int32_t bin_stride, int64_t R, int32_t S, int32_t contraction,
                     const float* aabb, const float* table, const float* scalings, int32_t L, int32_t log2_T,
                     const float* W1, const float* b1, const float* W2, const float* b2, int32_t H,
                     const float* ddensity, int32_t dedup_levels, float* dtable, float* dW1, float* db1,
                     float* dW2, float* db2, float* dfeat_out, void* stream);

/* N1 — per-ray training losses and their gradients in one ray-major pass (one warp per ray).
 * replaces: losses.interlevel_loss / distortion_loss, MSELoss and CrossEntropyLoss as called at
 *           nerfacto.py:468,483-489,505-515 and semantic_nerfw.py:253-279.
 * final level: sdist [R,S+1] spacing-space bin edges, w [R,S]; dw [R,S] receives d(distortion)/dw (NULL skips it).
 * proposal levels (HOST arrays of n_prop entries): S_prop, sdist_prop [R,Sp+1], w_prop [R,Sp], dw_prop [R,Sp]
 * receives d(interlevel)/dw_prop.  rgb/image/drgb [R,3] (rgb NULL skips the MSE); logits/dlogits [R,C], labels [R]
 * int64 (logits NULL skips the cross-entropy; dlogits = softmax - onehot, zero rows for ignore_index: the caller
 * divides by the count).  out [5], ACCUMULATED into: mean squared colour error, interlevel loss (sum over levels
 * of the [R,S] means), distortion loss (mean over rays), cross-entropy SUM, number of counted labels. */
int nkbs_ray_losses(int64_t R, int32_t S, const float* sdist, const float* w, float* dw, int32_t n_prop,
                    const int32_t* S_prop, const float* const* sdist_prop, const float* const* w_prop,
                    float* const* dw_prop, const float* rgb, const float* image, float* drgb, const float* logits,
                    const int64_t* labels, int32_t C, int64_t ignore_index, float* dlogits, float* out,
                    void* stream);

/* N1' — the same pass for a caller that owns the total loss (the step driver): term_weights = HOST [3] multipliers of
 * (colour MSE, interlevel, distortion) as in the loss dict of nerfacto.py:483-489 (1, interlevel_loss_mult,
 * distortion_loss_mult).  The gradients dw / dw_prop / drgb come out already multiplied by their term's weight, so they
 * seed the backward pass as they are; out [5] still receives the UNWEIGHTED terms; total [1] (optional, accumulated
 * into) receives the weighted sum of the three terms.  The cross-entropy outputs are as in nkbs_ray_losses.
 * scratch (may be NULL): DEVICE float [8] that is all zero between launches.  With it the CTAs accumulate into the
 * scratch, and the last CTA to finish WRITES out[0..4], total[0] and the weighted terms total[1..3] (total is then
 * [4]) and zeroes the scratch again — out / total need no zero fill before the launch.  Launches that share a
 * scratch must not overlap on the GPU. */
int nkbs_ray_losses_weighted(int64_t R, int32_t S, const float* sdist, const float* w, float* dw, int32_t n_prop,
                             const int32_t* S_prop, const float* const* sdist_prop, const float* const* w_prop,
                             float* const* dw_prop, const float* rgb, const float* image, float* drgb,
                             const float* logits, const int64_t* labels, int32_t C, int64_t ignore_index,
                             float* dlogits, float* out, const float* term_weights, float* total, float* scratch,
                             void* stream);

/* N3 — pinhole ray generation: ray_indices [N,3] int64 (camera, row, col), c2w [num_cameras,3,4], intrinsics
 * [num_cameras,4] = fx, fy, cx, cy.  Pixel centres at +0.5; camera looks down -z with +y up; directions are unit
 * vectors, directions_norm their length before normalisation, pixel_area the product of the distances to the rays
 * one pixel right and one pixel down (all three optional outputs may be NULL).
 * replaces: nerfstudio RayGenerator / Cameras.generate_rays as driven by dataparser/datamanager_z.py:96,252-278. */
int nkbs_generate_rays(int64_t N, const int64_t* ray_indices, const float* c2w, const float* intrinsics,
                       int64_t num_cameras, float* origins, float* directions, float* directions_norm,
                       float* pixel_area, int64_t* camera_indices, void* stream);

/* N3 — masked pixel sampling: keeps, in order, the candidates (image_indices[m], pixel_indices[m]) whose byte in
 * masks [num_images, pixels_per_image] is non-zero, up to R of them; if fewer survive they repeat cyclically.
 * *n_kept (device int32) receives the number of survivors.  block_ws NULL: one CTA walks the list (per-step candidate
 * lists, M <= 2^22).  block_ws = DEVICE int32 [ceil(M / 1024) + 1]: per-CTA counts + scan + ordered scatter, for
 * candidate lists of any length (the reference loads pixel chunks of 12.8 M, dataparser/datamanager_z.py:49).
 * replaces: dataparser/dataset_z.py:180-191. */
int nkbs_compact_masked(int64_t M, const int64_t* image_indices, const int64_t* pixel_indices, const uint8_t* masks,
                        int64_t pixels_per_image, int64_t R, int64_t* out_image, int64_t* out_pixel, int32_t* n_kept,
                        int32_t* block_ws, void* stream);

/* N3/N4 — semantic palette.  palette: device int32 [1 + 256] = {n, r|g<<8|b<<16 per class}.
 * color_to_label: rgb uint8 [N,3] -> labels int64 [N] (default_label when the colour is not in the palette);
 * replaces the per-pixel color2label loop of nerfacto.py:508-512.
 * semantic_colormap: logits [N,C] -> argmax labels (int64, may be NULL) and palette colours uint8 [N,3] (may be
 * NULL); replaces argmax(softmax) + colormap lookup of nerfacto.py:574-578 / semantic_nerfw.py:346-351. */
int nkbs_color_to_label(int64_t N, const uint8_t* rgb, const void* palette, int64_t default_label, int64_t* labels,
                        void* stream);
int nkbs_semantic_colormap(int64_t N, int32_t C, const float* logits, const void* palette, int64_t* labels,
                           uint8_t* rgb, void* stream);

/* K3+K4 — head MLPs on the tensor cores with their input assembled inside the kernel: the row
 * [SH16((d+1)/2) | geo15 | appearance32] (colour head) or [geo15 | transient16] (transient head) is built per
 * 128-sample tile straight into the first layer's operand tile, and the backward splits dX_0 into the geometry
 * gradient rows (dgeo, row stride dgeo_stride floats; may be NULL) and the per-image embedding gradient (demb
 * [num_images, emb_dim], accumulated; may be NULL), so neither the assembled [N,63] tensor nor its gradient ever
 * exists in HBM.  geo rows may be strided (geo_stride floats): the base MLP's [N,16] output is read in place.
 * replaces: nkbs_assemble_fwd/bwd + nkbs_mlp_tc_fwd/bwd for NerfactoField.get_outputs (nerfacto.py:398;
 *           semantic_nerfw.py:188).  Other arguments as nkbs_assemble_fwd and nkbs_mlp_tc_fwd/bwd. */
int nkbs_head_tc_fwd(int32_t sh_dim, int32_t geo_dim, int32_t emb_dim, const float* geo, int32_t geo_stride,
                     const float* directions, int32_t dir_per_ray, const int64_t* cam, int32_t cam_per_ray,
                     const float* emb, const float* emb_const, int32_t S, int64_t N, int32_t nl, const int32_t* dims,
                     const float* const* W, const float* const* b, int32_t relu_mask, int32_t out_act, int32_t fmt,
                     float* y, void* stream);
int nkbs_head_tc_bwd(int32_t sh_dim, int32_t geo_dim, int32_t emb_dim, const float* geo, int32_t geo_stride,
                     const float* directions, int32_t dir_per_ray, const int64_t* cam, int32_t cam_per_ray,
                     const float* emb, const float* emb_const, int32_t S, const float* dy, int64_t N, int32_t nl,
                     const int32_t* dims, const float* const* W, const float* const* b, int32_t relu_mask,
                     int32_t out_act, int32_t fmt, float* dgeo, int32_t dgeo_stride, int32_t dgeo_lead, float* demb,
                     float* const* dW, float* const* db, void* stream);
/* dgeo_lead: this many floats in FRONT of every dgeo row are set to zero by the kernel (dgeo pointing at column 1 of a
 * [N,16] gradient of the base MLP's output: dgeo_lead = 1 zeroes the density column, no fill kernel needed). */

/* K4 — head-MLP input assembly  h = [SH16((d+1)/2) | geo | embedding row]  and its backward.
 * replaces: SHEncoding + Embedding + torch.cat in NerfactoField.get_outputs (nerfacto.py:398;
 *           semantic_nerfw.py:188).  cam NULL -> emb_const [emb_dim] is used (eval). */
int nkbs_assemble_fwd(int64_t N, int32_t S, int32_t sh_dim, int32_t geo_dim, int32_t emb_dim,
                      const float* directions, int32_t dir_per_ray, const float* geo, const int64_t* cam,
                      int32_t cam_per_ray, const float* emb, const float* emb_const, int64_t num_images,
                      float* h, void* stream);
int nkbs_assemble_bwd(int64_t N, int32_t S, int32_t sh_dim, int32_t geo_dim, int32_t emb_dim,
                      const float* dh, float* dgeo, const int64_t* cam, int32_t cam_per_ray, float* demb,
                      void* stream);

/* K5 — samplers.  replaces: UniformLinDispPiecewiseSampler and PDFSampler inside
 * ProposalNetworkSampler (nerfacto.py:325-332,397; semantic_nerfw.py:132-137,187).
 * lin = torch.linspace(0,1,S+1); u_base = torch.linspace(0,1-1/nb,nb) (device copies of the host
 * tables).  t_rand/rand NULL = eval (no jitter / centred).  anneal_dev: optional device scalar that
 * overrides `anneal` (lets a captured CUDA graph see a new value every replay).  out_inds (int64 [R,nb]) may be NULL. */
int nkbs_initial_sampler(int64_t R, int32_t S, const float* lin, const float* t_rand, int32_t rand_per_bin,
                         const float* nears, const float* fars, float* sbins, float* ebins, void* stream);
int nkbs_pdf_sample(int64_t R, int32_t S, int32_t num_bins, const float* weights, float anneal,
                    const float* anneal_dev,
                    float histogram_padding, float eps, const float* sbins, const float* u_base,
                    const float* rand, int32_t rand_per_bin, const float* nears, const float* fars,
                    float* out_sbins, float* out_ebins, int64_t* out_inds, void* stream);

/* K6/K7 — ray-major compositing.  replaces: RaySamples.get_weights + RGB/Accumulation/Depth/
 * Semantic/Uncertainty renderers (nerfacto.py:399-449; semantic_nerfw.py:190-229).
 * Exactly one of density / weights_in.  Every output may be NULL.  bg_mode 0 last_sample, 1 black,
 * 2 white.  depth_exp = clip(depth_exp_raw, steps.min(), steps.max()) over the whole batch;
 * depth_mask = 1 where the clip is the identity; minmax_ws = int32[2] workspace. */
int nkbs_composite_fwd(int64_t R, int32_t S, const float* density, const float* weights_in,
                       const float* starts, const float* ends, int32_t bin_stride, const float* rgb, const float* sem, int32_t C,
                       int32_t bg_mode, int32_t eval_mode, float* weights, float* rgb_out, float* acc_out,
                       float* depth_exp_raw, float* depth_exp, float* depth_mask, float* depth_med,
                       float* sem_out, int32_t* minmax_ws, void* stream);
int nkbs_composite_bwd(int64_t R, int32_t S, const float* density, const float* weights, const float* starts,
                       const float* ends, int32_t bin_stride, const float* rgb, const float* sem, int32_t C, int32_t bg_mode,
                       int32_t sem_weight_grad, const float* acc, const float* depth_raw,
                       const float* g_weights, const float* g_rgb, const float* g_acc, const float* g_depth,
                       const float* g_sem, float* d_density, float* d_weights, float* d_rgb, float* d_sem,
                       void* stream);

/* K6 + K5 — one proposal level's RaySamples.get_weights (+ median depth, may be NULL) and the NEXT level's PDFSampler in
 * one ray-major launch: the warp that has computed a ray's weights resamples that ray straight from shared memory (no
 * [R,S] weights round trip, one launch instead of two).  density [R,S]; ebins / sbins [R,S+1] euclidean / spacing bin
 * edges of this level; the remaining arguments as nkbs_pdf_sample (rand: [R] per-ray jitter or NULL).  Same arithmetic
 * as nkbs_composite_fwd followed by nkbs_pdf_sample (bit-identical outputs).
 * replaces: RaySamples.get_weights + PDFSampler inside ProposalNetworkSampler (nerfacto.py:325-332,397). */
int nkbs_weights_resample_fwd(int64_t R, int32_t S, const float* density, const float* ebins, const float* sbins,
                              const float* nears, const float* fars, int32_t num_bins, float anneal,
                              const float* anneal_dev, float histogram_padding, float eps, const float* u_base,
                              const float* rand, float* weights, float* depth_med, float* out_sbins, float* out_ebins,
                              void* stream);

/* N2 — multi-tensor Adam (torch.optim.Adam semantics, amsgrad off).
 * replaces: the Adam(lr=1e-3, eps=1e-15) optimizers the reference configures for the "fields" and
 *           "proposal_networks" groups (train.py:85-98; groups nerfacto.py:359-363).
 * params/grads/exp_avg/exp_avg_sq: HOST arrays of num_tensors DEVICE pointers; numel: HOST.
 * step_dev: device scalar, the step count t >= 1 as float (bias corrections 1-beta^t are computed on the
 * device so a captured CUDA graph sees a new t every replay).  inv_scale_dev: optional device scalar
 * multiplied into every gradient (1/world after a sum-allreduce; a GradScaler's unscale).  zero_grads != 0: every
 * gradient element is overwritten with 0 after it has been read, so the backward kernels of the next step can
 * accumulate into the same buffers without a separate fill pass. */
int nkbs_adam_step(int32_t num_tensors, float* const* params, float* const* grads,
                   float* const* exp_avg, float* const* exp_avg_sq, const int64_t* numel, float lr,
                   float beta1, float beta2, float eps, float weight_decay, float* step_dev,
                   const float* inv_scale_dev, int32_t zero_grads, uint32_t* step_counter, void* stream);
/* step_counter NULL: *step_dev is the count t of THIS step (the caller advanced it).  step_counter = a device uint32
 * that is 0 between launches: *step_dev is the count before this step, the kernel uses t = *step_dev + 1 and stores it
 * when its last CTA finishes (single launch: at most 24 tensors), so no separate increment kernel is needed. */

/* N2 (data parallel) — reduce-scatter + Adam + all-gather in one kernel over NVLink peer memory.  grad_peers /
 * param_peers: HOST arrays of `world` DEVICE pointers to the flat gradient / parameter buffer of every rank
 * (symmetric memory: same layout on every GPU, peer-mapped).  This rank owns elements [lo, lo+n): it averages the
 * peers' gradients of that slice, updates its own shard of the moments (exp_avg, exp_avg_sq: [n]) and stores the new
 * parameters into every rank's buffer.  The caller orders it between two cross-rank barriers.
 * replaces: NCCL allreduce of the gradients + nkbs_adam_step (datamanager_z.py:225-231 DDP setup; train.py:85-98). */
int nkbs_dp_adam_step(int32_t world, int32_t rank, const float* const* grad_peers, float* const* param_peers,
                      int64_t lo, int64_t n, float* exp_avg, float* exp_avg_sq, float lr, float beta1, float beta2,
                      float eps, float weight_decay, float* step_dev, uint32_t* step_counter,
                      const float* grad_multicast, float* param_multicast, void* stream);
/* grad_multicast / param_multicast (both or neither): the NVSwitch multicast (NVLS) addresses of the two symmetric
 * buffers.  With them the gradient sum of the owned slice is one `multimem.ld_reduce` (reduced inside the switch) and
 * the new parameters one `multimem.st` (replicated by the switch) per 16 bytes, instead of `world` peer loads and
 * `world` peer stores. */

/* nkbs_density_bwd, dfeat_out (may be NULL): DEVICE float [N, 2L].  When given, the gradient w.r.t. the hash-grid
 * features of every sample is written there INSTEAD of being scattered into dtable (which may then be NULL); the caller
 * follows with nkbs_encode_bwd_rays.  The weight / bias gradients are accumulated as usual. */

/* K2 (ray-major) — gradient scatter for samples that come as rays (origins / directions [R,3], starts / ends [R,S] with
 * row stride bin_stride, S <= 256): one warp per ray, every lane walks a contiguous run of samples and sums the corner
 * gradients of consecutive samples that share a cell in registers; the table sees one set of reductions per cell
 * crossing.  dfeat [R*S, 2L] must be zero for samples outside the grid; dtable [L*T, 2] is ACCUMULATED into.
 * replaces: the autograd of HashEncoding inside HashMLPDensityField (nerfacto.py:299-311) for the proposal levels. */
int nkbs_encode_bwd_rays(int64_t R, int32_t S, const float* origins, const float* directions, const float* starts,
                         const float* ends, int32_t bin_stride, int32_t contraction, const float* aabb,
                         const float* dfeat, const float* scalings, int32_t L, int32_t log2_T, float* dtable,
                         void* stream);

/* One proposal level of ProposalNetworkSampler in one ray-major pass (forward): the level's bins (gen_bins != 0:
 * produced here by the initial piecewise sampler from lin [S+1] and the per-ray jitter t_rand [R] (NULL in
 * evaluation) and WRITTEN to sbins / ebins [R,S+1]; gen_bins == 0: READ from them) -> density [R,S] of the fused
 * HashMLPDensityField (arguments as nkbs_density_fwd) -> weights [R,S] (RaySamples.get_weights) and the median depth
 * [R] (may be NULL) -> the next level's num_bins spacing / euclidean bin edges out_sbins / out_ebins [R,num_bins] by
 * inverse-CDF resampling (arguments as nkbs_pdf_sample; rand [R] per-ray jitter, NULL = centred evaluation points).
 * Same arithmetic as the four stand-alone entry points it fuses; nothing else leaves the SM.
 * replaces: one iteration of ProposalNetworkSampler.generate_ray_samples (nerfacto.py:325-332,397;
 *           semantic_nerfw.py:132-137,187). */
int nkbs_proposal_level_fwd(int64_t R, int32_t S, const float* origins, const float* directions, const float* nears,
                            const float* fars, int32_t gen_bins, const float* lin, const float* t_rand, float* sbins,
                            float* ebins, int32_t contraction, const float* aabb, const float* table,
                            const float* scalings, int32_t L, int32_t log2_T, const float* W1, const float* b1,
                            const float* W2, const float* b2, int32_t H, float* density, float* weights,
                            float* depth_med, int32_t num_bins, float anneal, const float* anneal_dev,
                            float histogram_padding, float eps, const float* u_base, const float* rand,
                            float* out_sbins, float* out_ebins, void* stream);

/* N4 — scalar colour maps of the rendered frame.  lut: DEVICE float [256,3] colour table (the caller's data:
 * matplotlib's "viridis" / "turbo" tables in nerfstudio).  apply_colormap: image [N] in [0,1] -> rgb [N,3] =
 * lut[(long)(image * 255)].  apply_depth_colormap: depth [N] normalised by (near, far) — each either given
 * (have_* != 0) or the minimum / maximum of `depth` itself, reduced on the device into minmax_ws (2 x int32) —
 * as clip((d - near) / (far - near + 1e-10), 0, 1), mapped through lut, and with accumulation [N] (may be NULL)
 * blended over white: colour * acc + (1 - acc).
 * replaces: nerfstudio utils.colormaps.apply_colormap / apply_depth_colormap (nerfacto.py:574-578,608-613;
 *           semantic_nerfw.py:305). */
int nkbs_apply_colormap(int64_t N, const float* image, const float* lut, float* rgb, void* stream);
int nkbs_apply_depth_colormap(int64_t N, const float* depth, const float* accumulation, const float* lut,
                              int32_t have_near, float near_plane, int32_t have_far, float far_plane,
                              int32_t* minmax_ws, float* rgb, void* stream);

/* Expected depth (DepthRenderer("expected")) of R rays with NO allocation by the library or the caller beyond inputs
 * and output: the scratch that nkbs_composite_fwd takes from the caller (unclipped depth per ray, the batch-wide
 * min / max sample position) lives in a device array built into the library (4 MiB + 256 B per GPU that loads it, held
 * until exit).  weights [R,S], starts / ends [R,S] with row stride bin_stride (0 = S), depth out [R].
 * R > NKBS_NOALLOC_MAX_RAYS is refused (-1, nkbs_last_error says so).  Calls no cudaMalloc / cudaFree; capturable.
 * All calls share the scratch: calls that overlap on the GPU (different streams, concurrent graphs) race on it;
 * calls ordered on one stream are safe (every call re-initialises what it reads).
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it). */
#define NKBS_NOALLOC_MAX_RAYS (1 << 20)
int nkbs_expected_depth_noalloc(int64_t R, int32_t S, const float* weights, const float* starts, const float* ends,
                                int32_t bin_stride, float* depth, void* stream);

/* Zero `bytes` bytes of device memory on `stream` (cudaMemsetAsync: a memset node inside a captured CUDA graph).
 * Used for gradient buffers that cannot be re-zeroed by the optimizer kernel that reads them (peers still read
 * them over NVLink until the closing barrier of nkbs_dp_adam_step).
 * replaces: optimizer.zero_grad() of the nerfstudio Trainer loop (train.py:85-98 optimizers). */
int nkbs_zero_async(void* dev_ptr, int64_t bytes, void* stream);

/* Deterministic gradient accumulation (opt-in).  nkbs_fixed_point_register pairs a DEVICE float gradient buffer
 * grad_base [n] with a zero-filled DEVICE int64 shadow [n] (shadow == NULL removes the pair).  From then on every
 * backward entry point (nkbs_encode_bwd, nkbs_encode_bwd_rays, nkbs_density_bwd, nkbs_mlp_bwd, nkbs_mlp_tc_bwd,
 * nkbs_head_tc_bwd, nkbs_assemble_bwd) whose parameter-gradient destinations lie inside grad_base adds its partial sums to
 * the shadow as 64-bit fixed point (2^log2_scale units per 1.0; log2_scale in [16, 60]; |sum| must stay below
 * 2^(63 - log2_scale), partial sums below 2^-log2_scale are dropped) instead of issuing float atomics: integer additions
 * commute, so the result does not depend on the arrival order of warps and CTAs and two runs of the same step produce
 * the same bits.  nkbs_fixed_point_flush(dst, n): dst[i] += shadow[i] / 2^log2_scale and shadow[i] = 0 for a range
 * [dst, dst + n) inside a registered buffer — one thread per element, no atomics; call it before the optimizer reads the
 * gradients (dst an even number of elements into the buffer, shadow 16-byte aligned).  The registry is process-wide host state read at launch time (so a captured graph keeps the mode it was
 * captured in); registering is not thread-safe against concurrent launches into the same buffer.
 * replaces: nothing in the reference (torch's index_put / scatter backward there is non-deterministic as well); answers
 *           north_star's "atomic-free gradient scatter" with an order-independent one. */
int nkbs_fixed_point_register(float* grad_base, int64_t n, int64_t* shadow, int32_t log2_scale);
int nkbs_fixed_point_flush(float* dst, int64_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif
