/*
 * mdistiller_b200 -- C-ABI of the B200 (sm_100a) distillation tail.
 *
 * One shared library, plain C linkage, raw device pointers + sizes + an explicit
 * stream.  No entry point below allocates, frees or retains device memory: every
 * buffer, the reduction workspace included, belongs to the caller and must stay
 * alive until the work queued on `stream` has finished (one variant,
 * mdb200_distill_loss_static_scratch, uses a fixed array inside the library instead
 * of a caller workspace).  Calls return 0 on
 * success, a negative MDB200_E* code for a rejected argument, or a positive
 * cudaError_t when the launch failed; nothing is thrown across the boundary.
 *
 * The reference (youngyoonii/mdistiller-ddp) has no native layer; each entry point
 * cites the Python it replaces (paths relative to the reference root).
 */
#ifndef MDISTILLER_B200_H_
#define MDISTILLER_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MDB200_ABI_VERSION 3

#if defined(__GNUC__)
#define MDB200_API __attribute__((visibility("default")))
#else
#define MDB200_API
#endif

/* argument errors (negative so they never collide with cudaError_t) */
#define MDB200_EINVAL   (-1)   /* null pointer, non-positive size, bad mode/flag      */
#define MDB200_ESHAPE   (-2)   /* leading dimension smaller than the row length      */
#define MDB200_EWORKSPC (-3)   /* workspace missing or smaller than required         */
#define MDB200_EALIGN   (-4)   /* pointer not 4-byte aligned                         */
#define MDB200_ESCRATCH (-5)   /* launch needs more CTA slots than the built-in scratch */

/* loss kinds */
#define MDB200_LOSS_DKD 0      /* alpha*TCKD + beta*NCKD   (distillers/DKD.py:8-32)  */
#define MDB200_LOSS_KD  1      /* T^2 * KL(p_t || p_s)     (distillers/KD.py:8-13)   */

/* cross-entropy handling of mdb200_distill_loss */
#define MDB200_CE_NONE     0   /* distillation term only                              */
#define MDB200_CE_SPLIT    1   /* CE gradient written to its own buffer (DOT trainer) */
#define MDB200_CE_COMBINED 2   /* CE gradient added into grad_kd (base trainer)       */

/* membership of a parameter range in DOT's two gradient lists (engine/dot.py:36-39,49-54) */
#define MDB200_DOT_BOTH      0 /* has task and KD gradient: m_task / m_kd             */
#define MDB200_DOT_TASK_ONLY 1 /* task gradient only: coefficient (m_task+m_kd)/2     */
#define MDB200_DOT_KD_ONLY   2 /* KD gradient only: weight decay joins the KD branch  */

MDB200_API int mdb200_abi_version(void);

/* Human-readable text for a code returned by any entry point (static storage). */
MDB200_API const char *mdb200_error_string(int code);

/* Bytes of device workspace mdb200_distill_loss needs (a constant; 256-byte
 * aligned buffer).  The first 64 bytes must be zero before the FIRST call; the
 * kernel leaves them zero again, so one memset at allocation time is enough.
 * Concurrent launches on different streams need different workspaces. */
MDB200_API size_t mdb200_loss_workspace_bytes(void);

/*
 * Fused logit loss: value AND gradient w.r.t. the student logits in one pass.
 *
 * Replaces dkd_loss (+ _get_gt_mask/_get_other_mask/cat_mask) at
 * mdistiller/distillers/DKD.py:8-51, kd_loss at mdistiller/distillers/KD.py:8-13,
 * the F.cross_entropy term at DKD.py:71 / KD.py:31 and the autograd backward of all
 * three.
 *
 *   z_student, z_teacher  [rows, classes] fp32, row stride ld_s / ld_t elements
 *   target                [rows] int64 class indices in [0, classes)
 *   kind                  MDB200_LOSS_DKD | MDB200_LOSS_KD (alpha, beta ignored for KD)
 *   kd_scale              outer factor on the distillation term: warm-up
 *                         min(epoch/WARMUP,1) (DKD.py:72) or KD_WEIGHT (KD.py:32)
 *   ce_mode, ce_scale     CE_WEIGHT * F.cross_entropy(z_student, target) handling
 *   grad_kd               [rows, classes] fp32 out, row stride ld_g:
 *                         d(kd_scale*loss_kd)/dz_student (+ CE part if CE_COMBINED)
 *   grad_ce               same shape/stride, written only for CE_SPLIT (else may be NULL)
 *   loss_out              2 floats on the device: {kd_scale*loss_kd, ce_scale*loss_ce}
 *                         (second is 0 for CE_NONE)
 *   workspace             see mdb200_loss_workspace_bytes
 *
 * The means are over `rows` (the per-rank batch), exactly like the reference.
 * Rows are reduced in a fixed order: results are bit-reproducible run to run.
 */
MDB200_API int mdb200_distill_loss(const float *z_student, const float *z_teacher, const int64_t *target,
                        int64_t rows, int64_t classes, int64_t ld_s, int64_t ld_t,
                        int kind, float alpha, float beta, float temperature,
                        float kd_scale, int ce_mode, float ce_scale,
                        float *grad_kd, float *grad_ce, int64_t ld_g,
                        float *loss_out, void *workspace, size_t workspace_bytes,
                        void *stream);

/*
 * mdb200_distill_loss without a caller workspace.  (No counterpart in the reference: it
 * exists for hosts of this library that cannot allocate device memory.)
 *
 * Same arguments, kernels and results as mdb200_distill_loss, minus `workspace`: the
 * partial-sum scratch is a fixed array inside the library (64 + 2 * 8 *
 * MDB200_STATIC_SCRATCH_CTAS bytes = 16 448 bytes, resident on every GPU that loads the
 * library until the process exits).  The call itself allocates and frees nothing, on the
 * host or the device, and uses no allocating library routine; inputs and outputs are
 * still the caller's buffers.
 *
 * Limit: the launch may use at most MDB200_STATIC_SCRATCH_CTAS thread blocks.  The grid
 * is chosen exactly as in mdb200_distill_loss: min(ceil(rows / R), SMs x resident blocks
 * per SM).  On a 148-SM B200 the register-resident kernels (classes and strides multiples
 * of 4, 16-byte aligned pointers, classes <= 2048) hold at most 4 blocks per SM, i.e. at
 * most 592 blocks, and always fit.  Shapes outside that path run one row per block with up
 * to 8 resident blocks per SM (KD) or 4 (DKD): those fit up to 1 024 rows in KD mode and
 * always in DKD mode.  Anything larger is refused with MDB200_ESCRATCH and nothing is
 * launched; use mdb200_distill_loss there.
 *
 * Sharing: all calls on a device use the same array.  The entry point clears the part
 * that must start at zero on `stream` before the kernel, so back-to-back calls on one
 * stream are fine; two calls whose kernels can overlap on the GPU (different streams,
 * no dependency) race on it and must be serialised by the caller.
 *
 * First call on each device: resolves the device address of the array
 * (cudaGetSymbolAddress with that device current) and caches it; make that first call
 * outside any stream capture.
 */
#define MDB200_STATIC_SCRATCH_CTAS 1024
MDB200_API int mdb200_distill_loss_static_scratch(const float *z_student, const float *z_teacher,
                        const int64_t *target, int64_t rows, int64_t classes, int64_t ld_s, int64_t ld_t,
                        int kind, float alpha, float beta, float temperature,
                        float kd_scale, int ce_mode, float ce_scale,
                        float *grad_kd, float *grad_ce, int64_t ld_g,
                        float *loss_out, void *stream);

/*
 * out[i] = grad[i] * (*upstream)   -- the autograd "grad_output" multiply
 * (what torch does behind loss.backward() for DKD.py:72 / KD.py:32), with the
 * scalar read on the device so the host never synchronises.  out may alias grad.
 */
MDB200_API int mdb200_scale_by_device_scalar(float *out, const float *grad, const float *upstream,
                                  int64_t n, void *stream);

/*
 * Distillation-Oriented Trainer update on flat fp32 ranges, in place.
 *
 * Replaces dot() at mdistiller/engine/dot.py:15-55 as driven by
 * DistillationOrientedTrainer.step_kd/step (dot.py:96-174): per element
 *   d   = grad_scale*g_task + wd*p;  v_task = first_task ? d : m_task*v_task + (1-damp)*d
 *   p  -= lr*v_task                  (v_task bypassed when momentum == 0, dot.py:31,40)
 *   v_kd = first_kd ? gk : m_kd*v_kd + (1-damp)*gk,  gk = grad_scale*g_kd
 *   p  -= lr*v_kd
 * `membership` selects the reference's coefficient/decay rules for ranges that are
 * in only one gradient list; the buffers of the absent branch may be NULL.
 * grad_scale = 1/world_size after a sum reduce-scatter, 1 otherwise.
 * 32 algorithmic bytes per element (BOTH), n elements, any alignment >= 4 bytes
 * (16-byte aligned pointers take the vector path).
 */
MDB200_API int mdb200_dot_step(float *p, const float *g_task, const float *g_kd,
                    float *v_task, float *v_kd, int64_t n,
                    float lr, float momentum, float momentum_kd, float weight_decay,
                    float dampening, float grad_scale, int first_task, int first_kd,
                    int membership, void *stream);

/*
 * Data-parallel DOT step as ONE kernel over NVLink peer memory: gradient sum over ranks ->
 * dual-momentum update -> parameters written to every rank, for the element range
 * [shard_lo, shard_lo + shard_len) this rank owns.
 *
 * Replaces DDP's bucketed all-reduce (tools/train.py:86) followed by dot()
 * (mdistiller/engine/dot.py:15-55) on this path, and the NCCL reduce-scatter / all-gather
 * pair of the unfused step.  Arithmetic is mdb200_dot_step's MDB200_DOT_BOTH case.
 *
 *   p_local / g_task_local / g_kd_local   this rank's FULL flat buffers (symmetric allocations:
 *                                         same size on every rank, mapped into every peer)
 *   *_multicast                           NVSwitch multicast address of the same buffer, or NULL;
 *                                         when given, sums use multimem.ld_reduce and the
 *                                         parameter write-back multimem.st (in-switch, "NVLS")
 *   *_peers                               host array[world] of every rank's base address of the
 *                                         buffer (own rank included); used when no multicast
 *   kd_is_reduced                         non-zero: g_kd_local[shard] already holds the SUM over
 *                                         ranks (reduce-scattered earlier, e.g. overlapped with
 *                                         the second backward); only the task gradient is fetched
 *   v_task, v_kd                          rank-local momentum, shard_len elements
 *   grad_scale                            normally 1/world
 *   sync_pads, rank, epoch                cross-rank ordering inside the kernel: host array[world] of every
 *                                         rank's SYNC PAD (64 uint32 of symmetric memory, zero before the first
 *                                         use; own rank included), this rank's index, and a launch counter that
 *                                         starts at 1 and grows by 1 with every launch that uses the pads (the
 *                                         same value on every rank) -- or 0, and the kernel keeps that count
 *                                         itself in the pad (the form that can be replayed from a CUDA graph;
 *                                         do not mix the two forms on one pad).  The kernel then waits until every rank's
 *                                         gradients are complete before its first fetch and does not finish
 *                                         until every rank has written this rank's copy of P.  One rank per GPU;
 *                                         a rank that never launches makes the others trap after 20 s.
 *                                         sync_pads == NULL: no handshake -- the caller must order the launch
 *                                         after every rank's gradients (a barrier before) and keep every rank
 *                                         from reading P until all kernels have finished (a barrier after).
 * shard_lo and shard_len are multiples of 4; world <= 16.
 */
MDB200_API int mdb200_dot_step_fused(float *p_local, float *p_multicast, float *const *p_peers,
                    const float *g_task_local, const float *g_task_multicast, float *const *g_task_peers,
                    const float *g_kd_local, const float *g_kd_multicast, float *const *g_kd_peers,
                    int kd_is_reduced, float *v_task, float *v_kd,
                    int64_t shard_lo, int64_t shard_len, int world,
                    float lr, float momentum, float momentum_kd, float weight_decay,
                    float dampening, float grad_scale, int first_task, int first_kd,
                    uint32_t *const *sync_pads, int rank, uint32_t epoch, void *stream);

/*
 * The same fused data-parallel step for the base trainer's single gradient bucket: gradient sum over ranks ->
 * momentum SGD (arithmetic of mdb200_sgd_step) on this rank's shard -> parameters written to every rank, one
 * kernel over NVLink peer memory.
 *
 * Replaces DDP's bucketed all-reduce (tools/train.py:86) followed by optim.SGD.step()
 * (mdistiller/engine/trainer.py:86-91, 321) and the NCCL reduce-scatter / all-gather pair of the unfused step.
 * Arguments as in mdb200_dot_step_fused (one bucket `g`, one momentum shard `v`; v may be NULL when momentum == 0).
 */
MDB200_API int mdb200_sgd_step_fused(float *p_local, float *p_multicast, float *const *p_peers,
                    const float *g_local, const float *g_multicast, float *const *g_peers, float *v,
                    int64_t shard_lo, int64_t shard_len, int world,
                    float lr, float momentum, float weight_decay, float dampening, float grad_scale,
                    int first_step, uint32_t *const *sync_pads, int rank, uint32_t epoch, void *stream);

/*
 * Multi-tensor gather: copy n_tensors per-parameter gradient tensors into one flat fp32 bucket in a
 * single launch (zero-filling the slices of parameters that have no gradient, src_ptrs[i] == NULL).
 *
 * Replaces the reference's reliance on `zero_grad(set_to_none=True)` + captured `p.grad` tensors
 * (mdistiller/engine/dot.py:111-123, trainer.py:427-431) and DDP's gradient-to-bucket copies
 * (tools/train.py:86) on the way into the flat step / reduce-scatter.
 *
 *   src_ptrs         host array[n_tensors] of device pointers (contiguous fp32 tensors) or NULL
 *   chunk_table_dev  device array[n_chunks][4] of int64 {tensor index, element count, source element
 *                    offset, bucket element offset}; shape-dependent only, built once by the caller;
 *                    chunks of one tensor must tile it exactly
 * 8 algorithmic bytes per gathered element.
 */
MDB200_API int mdb200_gather_flat(float *dst_flat, const float *const *src_ptrs, int n_tensors,
                    const int64_t *chunk_table_dev, int64_t n_chunks, void *stream);

/*
 * Metrics record of one iteration in one launch: record[5] = {*loss_ce, *loss_kd, top-1 hits,
 * top-5 hits, rows} for [rows, classes] fp32 logits and int64 labels.
 *
 * Replaces accuracy() (mdistiller/engine/utils.py:125-136: topk + eq + sums) and the per-scalar
 * packing on the metrics tail of train_iter (mdistiller/engine/trainer.py:277-290, 436-450); the
 * host code all-reduces the record once and reads it back once.  record[2] and record[3] must be
 * zero on entry (they are accumulated); loss pointers may be NULL (recorded as 0).  Equal logits
 * rank by class index (torch.topk leaves the order of equal values unspecified).
 */
MDB200_API int mdb200_metrics_record(const float *logits, const int64_t *target, int64_t rows, int64_t classes,
                    int64_t ld, const float *loss_ce, const float *loss_kd, float *record, void *stream);

/*
 * Single-momentum SGD on a flat range, in place: torch.optim.SGD(momentum,
 * weight_decay, dampening, nesterov=False) as built by BaseTrainer.init_optimizer
 * (mdistiller/engine/trainer.py:86-91) and stepped at trainer.py:321.
 * v may be NULL when momentum == 0.  grad_scale_dev (may be NULL) points at one float on
 * the device that also multiplies the gradient -- the clipping coefficient of
 * mdb200_clip_coefficient or the unscale factor of mdb200_amp_step_scale; a NEGATIVE value
 * means "skip": the kernel returns without touching p or v.  20 algorithmic bytes per element.
 */
MDB200_API int mdb200_sgd_step(float *p, const float *g, float *v, int64_t n,
                    float lr, float momentum, float weight_decay, float dampening,
                    float grad_scale, const float *grad_scale_dev, int first_step, void *stream);

/*
 * Adam / AdamW on a flat range, in place: torch.optim.Adam / AdamW (no amsgrad) as built by
 * BaseTrainer.init_optimizer (mdistiller/engine/trainer.py:92-107, SOLVER.TYPE "Adam" /
 * "AdamW", configs/imagenet/optim/*.yaml) and stepped at trainer.py:321.  `step` is the
 * 1-based step count (bias corrections 1 - beta^step are formed in double on the host side
 * of this call, like torch's Python floats).  decoupled != 0 selects AdamW's
 * p *= 1 - lr*wd; otherwise weight decay joins the gradient.  m, v: first and second moment.
 * The betas are doubles so that 1 - beta2 keeps the precision torch's Python floats give it.
 * 28 algorithmic bytes per element.
 */
MDB200_API int mdb200_adam_step(float *p, const float *g, float *m, float *v, int64_t n,
                    float lr, double beta1, double beta2, float eps, float weight_decay, int decoupled,
                    int64_t step, float grad_scale, const float *grad_scale_dev, void *stream);

/*
 * Gradient clipping by global norm, nn.utils.clip_grad_norm_ at mdistiller/engine/trainer.py:316-320,
 * in two device-side pieces so the host never reads the norm:
 *   mdb200_sum_squares      *out = sum over [0, n) of (scale * x[i])^2 in fp64, fixed summation
 *                           order; workspace >= 64 + 1024*8 bytes, first 64 bytes zero before the
 *                           first call (the loss workspace qualifies)
 *   (host code all-reduces *out over ranks when the gradient is sharded)
 *   mdb200_clip_coefficient *coef = min(1, max_norm / (sqrt(*total_sumsq) + 1e-6))
 * and the step kernels take `coef` as grad_scale_dev.
 */
MDB200_API int mdb200_sum_squares(const float *x, int64_t n, float scale, double *out,
                    void *workspace, size_t workspace_bytes, void *stream);
MDB200_API int mdb200_clip_coefficient(const double *total_sumsq, float max_norm, float *coef, void *stream);

/*
 * Mixed-precision loss scaling decided on the device: the AMP branch of BaseTrainer.backward_loss
 * (mdistiller/engine/trainer.py:304-313: GradScaler.scale / unscale_ / clip_grad_norm_ / step / update) without a
 * host read of the overflow flag and without a pass that rewrites the gradients.
 *   total_sumsq     sum of squares of the SCALED (and rank-averaged) gradient from mdb200_sum_squares, all-reduced
 *                   over ranks by the host code; non-finite exactly when some gradient element is inf or nan
 *   scale, growth_tracker   the scaler's state on the device, updated in place (GradScaler.update: x backoff_factor
 *                   after an overflow, x growth_factor after growth_interval clean steps in a row)
 *   max_norm        SOLVER.GRAD_CLIP on the unscaled gradient, 0 = no clipping
 *   coef            out: 1/scale (x clipping coefficient), to be passed as grad_scale_dev of mdb200_sgd_step /
 *                   mdb200_adam_step; NEGATIVE after an overflow, which those kernels take as "skip this update"
 */
MDB200_API int mdb200_amp_step_scale(const double *total_sumsq, float *scale, int *growth_tracker,
                    float growth_factor, float backoff_factor, int growth_interval, float max_norm,
                    float *coef, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MDISTILLER_B200_H_ */
