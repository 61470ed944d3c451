/*
 * yolohead.h -- C ABI of libyolohead.so, the B200 (sm_100a) implementation of the
 * YOLOv1 detection-head hot path of mattiaskvist/yolo-v1.
 *
 * The reference is pure Python; its "FFI" for this path is the Python call surface
 * of src/yolo/loss.py, src/yolo/metrics.py and src/yolo/inference.py.  Each entry
 * point below names the reference code it replaces (paths relative to the
 * reference checkout).  The Python host in yolo_v1_b200/ binds these with ctypes,
 * passing tensor.data_ptr() and torch.cuda.current_stream().cuda_stream.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - nothing is allocated or freed here: the caller owns all buffers, including
 *     the workspace whose size the matching *_workspace_bytes() call returns;
 *   - all work is enqueued on `stream` (a cudaStream_t); no call synchronises;
 *   - return value: 0 = ok, < 0 = YH_ERR_*, > 0 = a cudaError_t from a launch;
 *     yh_last_error() gives a message for the calling thread's last failure;
 *   - thread-safe for calls on distinct streams with distinct buffers.
 *   - layout constants: one image = S*S cells of D = B*5+C floats
 *     [x0,y0,w0,h0,c0, x1,...,c1, p0..pC-1]; candidate k = (row*S + col)*B + b.
 *     This build is specialised for S=7, B=2, C=20 (anything else returns
 *     YH_ERR_UNSUPPORTED; the Python host raises NotImplementedError).
 */
#ifndef YOLOHEAD_H
#define YOLOHEAD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YH_VERSION 100

#define YH_ERR_BAD_ARG (-1)
#define YH_ERR_UNSUPPORTED (-2)
#define YH_ERR_WORKSPACE (-3)
#define YH_ERR_ALIGNMENT (-4)

/* pred_dtype */
#define YH_F32 0
#define YH_F16 1
#define YH_BF16 2
#define YH_TF32 3 /* yh_head_linear_loss only: fp32 operands read by the tensor core as TF32, one pass (torch's allow_tf32) */

/* iou_mode: which of the reference's IoU definitions NMS uses */
#define YH_IOU_METRICS 0 /* metrics.py:298-341: inter/union, union == 0 -> 0            */
#define YH_IOU_INFER 1   /* inference.py:212-249: inter/(a1 + a2 - inter + 1e-6)        */

/* order of the kept list */
#define YH_ORDER_CLASS_GROUPED 0 /* metrics.py:277-296: class groups by first appearance */
#define YH_ORDER_GLOBAL_CONF 1   /* inference.py:302-317: global confidence order        */

#define YH_MAX_THRESHOLDS 13 /* 4 subsets * T + 1 overall bit + 8 class bits fit 64 */

int yh_version(void);
const char *yh_last_error(void);
/* kernels this library has launched in this process so far (all threads) */
long long yh_launch_count(void);

/* ---------------------------------------------------------------- loss ----
 * Replaces YOLOLoss.forward + compute_iou + their autograd backward
 * (src/yolo/loss.py:55-172, 174-212).
 * parts[5] (device, fp32) = {total, coord, conf_obj, conf_noobj, class}, each
 * already divided by N (loss.py:162-170).  grad_pred (same dtype and shape as
 * pred; NULL = forward only) receives d total / d pred for an upstream gradient
 * of 1.  One fused pass: 17,640 B of HBM traffic per image in fp32. */
size_t yh_loss_workspace_bytes(int N, int S, int B, int C);
int yh_loss_fwd_bwd(const void *pred, int pred_dtype, const float *tgt, int N, int S, int B, int C,
                    float lambda_coord, float lambda_noobj, float *parts, void *grad_pred,
                    void *workspace, size_t workspace_bytes, void *stream);

/* grad[i] *= *scale (device scalar; the autograd upstream gradient, e.g. the
 * GradScaler factor at training/trainer.py:75).  Every CTA returns at once when
 * *scale == 1, so the plain fp32 training path costs no extra traffic. */
int yh_scale_inplace(void *grad, int dtype, size_t n_elems, const float *scale, void *stream);

/* yh_loss_fwd_bwd without a caller-supplied workspace.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  The kernel's scratch (ticket word + per-CTA partial sums) is a 128 KiB
 * file-scope device array (yh_loss_static_scratch_bytes()), held by every GPU that loads the library until the
 * process exits.  Same arguments, results and error codes as yh_loss_fwd_bwd otherwise; a batch whose
 * yh_loss_workspace_bytes(N, S, B, C) exceeds the array is rejected with YH_ERR_WORKSPACE (S=7, B=2, C=20:
 * at most 10,678 images per call).  Allocates nothing, frees nothing, synchronises nothing.  All calls on a
 * device share the array: calls ordered on one stream are independent (what must start at zero is cleared on
 * the stream by every call), calls that OVERLAP on the GPU race on it and return wrong sums. */
size_t yh_loss_static_scratch_bytes(void);
int yh_loss_fwd_bwd_static(const void *pred, int pred_dtype, const float *tgt, int N, int S, int B, int C,
                           float lambda_coord, float lambda_noobj, float *parts, void *grad_pred,
                           void *stream);

/* ------------------------------------------------------- decode + NMS ----
 * Replaces mAPMetric._parse_predictions + _apply_nms (metrics.py:173-218,
 * 258-296) and YOLOInference.parse_predictions + non_max_suppression
 * (inference.py:141-210, 251-317) for a batch of N images, one warp per image.
 * Outputs, per image n (K = S*S*B slots each):
 *   n_keep[n]          number of kept detections
 *   keep_k[n*K + q]    candidate index of the q-th kept detection, in `order`
 *   keep_cls[n*K + q]  its class;   keep_conf[n*K + q]  its confidence (fp64)
 *   keep_box[(n*K + q)*4 ..]  (x_abs, y_abs, w, h) in fp64 as the reference builds them
 *   n_cand[n]          candidates above conf_thr before NMS (nullable)
 *   bad_range[n]       nonzero if any candidate above conf_thr has x,y,w,h or
 *                      confidence outside [0,1] -- where the reference's pydantic
 *                      schema raises (schemas.py:9-16,88) (nullable)
 * keep_cls / keep_conf / keep_box are nullable. */
int yh_decode_nms(const float *pred, int N, int S, int B, int C, double conf_thr, double nms_thr,
                  int iou_mode, int order, int32_t *n_keep, uint8_t *keep_k, uint8_t *keep_cls,
                  double *keep_conf, double *keep_box, int32_t *n_cand, uint8_t *bad_range,
                  void *stream);

/* ------------------------------------------- mAP update (fused, records) ----
 * Replaces mAPMetric.update (metrics.py:78-100) AND the per-detection part of
 * compute(): best-IoU ground truth, greedy TP/FP at every threshold for the four
 * ground-truth subsets (all / small / medium / large, metrics.py:343-413,
 * 444-491, 493-633).  Matching is image-local, so each image emits final flags.
 * Per image n, for its q-th kept detection in (class asc, confidence desc,
 * candidate asc) order:
 *   rec_key[n*K+q]   = ~orderable(conf): ascending key order = confidence descending
 *                      (orderable() maps a double to a uint64 with the same order)
 *   rec_flags[n*K+q] = class << 56
 *                      | bit s*T+t: TP for subset s at thresholds[t]
 *                      | bit 4*T:   TP in the overall pass at IoU 0.5
 *   rec_k[n*K+q]     = candidate index
 *   rec_box[(n*K+q)*24 B] = the box as the reference stores it: x_abs, y_abs as fp64 and
 *                      (w, h) as two fp32 in the third 8 bytes (nullable; feeds yh_list_form)
 *   n_keep[n], n_gt[n]; gt_count[s*C + c] += ground truths (atomic; zero it first)
 *   gt_cell / gt_cls [n*S*S+g], gt_raw[(n*S*S+g)*4..]: cell index, class and raw
 *                      (x, y, w, h) of every ground truth (all three nullable together)
 * thresholds_host: T doubles on the HOST (copied into the launch parameters). */
int yh_map_update(const float *pred, const float *tgt, int N, int S, int B, int C, double conf_thr,
                  double nms_thr, const double *thresholds_host, int T, int32_t *n_keep,
                  uint64_t *rec_key, uint64_t *rec_flags, uint8_t *rec_k, void *rec_box,
                  int32_t *n_gt, uint8_t *gt_cell, uint8_t *gt_cls, float *gt_raw,
                  int32_t *gt_count, void *stream);

/* yh_map_update with the loss riding on it (S=7, B=2, C=20, fp32 predictions): a validation or monitoring
 * step that wants YOLOLoss.forward (+ backward) AND mAPMetric.update of the same batch (trainer.py:87,174 next to
 * metrics.py:709) reads the two tensors once instead of twice.  While a prediction cell is staged for decoding its
 * loss terms are computed (loss.py:98-159, the arithmetic of yh_loss_fwd_bwd) and the cell is replaced by
 * d loss / d prediction, which leaves as the gradient tensor (grad_pred nullable = forward only).  parts[5] as
 * yh_loss_fwd_bwd (device, already / N; summed in a fixed order, so repeatable).  Records / counts as
 * yh_map_update, without its optional list outputs.  workspace: yh_map_update_loss_workspace_bytes(N),
 * zero-filled before the first use (the call leaves it ready for the next one), 256-byte aligned.
 * phases: 3 = everything on `stream`; 1 = the update kernel only (records, counts, gradient of the no-object
 * cells); 2 = the loss tail only (object cells -> their gradient cells, then parts[5]) -- a caller that wants the
 * record pipeline (compaction, exchange, sort) to start right behind the update kernel issues phase 2 with the same
 * arguments on another stream, ordered after phase 1 by an event. */
size_t yh_map_update_loss_workspace_bytes(int N);
int yh_map_update_loss(const float *pred, const float *tgt, int N, int S, int B, int C, double conf_thr,
                       double nms_thr, const double *thresholds_host, int T, int32_t *n_keep,
                       uint64_t *rec_key, uint64_t *rec_flags, uint8_t *rec_k, int32_t *n_gt,
                       int32_t *gt_count, float lambda_coord, float lambda_noobj, float *parts,
                       float *grad_pred, void *workspace, size_t workspace_bytes, int phases, void *stream);

/* The reference keeps its state as two Python lists (metrics.py:70-71, 99-100).
 * yh_list_form expands the records of yh_map_update into that form on demand:
 *   rec_pos[n*K+q]       position of record q in all_predictions[n] (class groups in
 *                        first-appearance order of the confidence-sorted list,
 *                        metrics.py:274-279)
 *   rec_box[(n*K+q)*4..] (x_abs, y_abs, w, h) fp64 as metrics.py:200-203 builds them
 *                        (expanded from the 24-byte rec_box3 records of yh_map_update)
 *   gt_box[(n*S*S+g)*4..] the same for ground truths (metrics.py:247-248) */
int yh_list_form(const int32_t *n_keep, const uint64_t *rec_key, const uint64_t *rec_flags,
                 const uint8_t *rec_k, const void *rec_box3, const int32_t *n_gt,
                 const uint8_t *gt_cell, const float *gt_raw, int N, int S, int B, int C,
                 uint8_t *rec_pos, double *rec_box, double *gt_box, void *stream);

/* Ground truths of a batch in the reference's list form (metrics.py:220-256):
 * gt_cls[n*S*S + g], gt_box[(n*S*S + g)*4] fp64, n_gt[n]. */
int yh_parse_gt(const float *tgt, int N, int S, int B, int C, int32_t *n_gt, uint8_t *gt_cls,
                double *gt_box, void *stream);

/* Compacts the fixed-stride records of yh_map_update into dense arrays in image
 * order.  running_total (device int64, nullable) chains calls: the records of this
 * batch are appended behind the *running_total records already in out_key /
 * out_flags and the scalar is advanced, all on the stream (no host round trip).
 * offsets[n] = position of image n's first record in the dense arrays,
 * offsets[N] = position one past the last.  Uses `workspace` of
 * yh_compact_workspace_bytes(N). */
size_t yh_compact_workspace_bytes(int N);
int yh_compact_records(const int32_t *n_keep, const uint64_t *rec_key, const uint64_t *rec_flags,
                       int N, int K, int64_t *offsets, uint64_t *out_key, uint64_t *out_flags,
                       int64_t *running_total, void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------------------ AP reduce ----
 * Replaces the sort / cumsum / 11-point interpolation of
 * mAPMetric._calculate_ap_for_class and _calculate_ap_for_size_class
 * (metrics.py:382, 415-436, 606, 635-649) over M compact records.
 * key/flags are scratch afterwards (the stable LSD radix sort ping-pongs between
 * them and the workspace; ties keep image order then per-image order -- the
 * reference's stable sorted()).  The class is read from flags >> 56.
 * The sort runs four 8-bit passes over the leading 32 bits of (class, varying key
 * bits) and then orders the few short runs of records that agree in those bits;
 * inputs this does not suit (long unordered runs, > 59 varying key bits) are
 * re-sorted by all digits inside the same call.
 *   ap[(s*C + c)*T + t]  fp64, bit-exact with the reference's arithmetic
 *   totals[0] = overall TP (bit 4*T), totals[1] = M  (int64)
 * gt_count[4*C] as produced by yh_map_update. */
size_t yh_ap_workspace_bytes(int64_t M, int C, int T);
/* M_dev (device int64, nullable): the actual record count, min(*M_dev, M); M is then
 * only the capacity the buffers and the workspace were sized for, so the host never
 * has to read the count back.  packed (nullable, 4*C*T + 3 + 4*C doubles) receives
 * everything a caller reads back in one buffer: ap[], totals[0], totals[1], an
 * internal error flag (0 = ok), gt_count[] as doubles. */
int yh_ap_reduce(uint64_t *key, uint64_t *flags, int64_t M, const int64_t *M_dev,
                 const int32_t *gt_count, int C, int T, double *ap, int64_t *totals, double *packed,
                 void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------- multi-GPU exchange ----
 * Average precision is independent per class (metrics.py:126-141, 382), so N ranks
 * can each sort and reduce a subset of the classes.  yh_partition_records lays the
 * M compact records of one rank out as n_parts contiguous runs (run p = records
 * whose class c has owner_of_class_host[c] == p), each run in the input (image)
 * order, and writes the run lengths to part_count[n_parts] (device, int64): the
 * split sizes of one all_to_all_single.  n_parts <= 32.  M_dev as in yh_ap_reduce:
 * nullable device count, M is then the capacity. */
size_t yh_partition_workspace_bytes(int64_t M);
int yh_partition_records(const uint64_t *key, const uint64_t *flags, int64_t M, const int64_t *M_dev,
                         const uint8_t *owner_of_class_host, int C, int n_parts, uint64_t *out_key,
                         uint64_t *out_flags, int64_t *part_count, void *workspace,
                         size_t workspace_bytes, void *stream);

/* -------------------------------- fused partition + exchange over peer memory ----
 * The same stable partition, with each run written straight into its owner's receive
 * buffer through peer-mapped memory (NVLink / NVSwitch): partition and all-to-all are one
 * kernel and the run lengths never visit the host.  There is no reference counterpart
 * (the reference is single-process); it serves DistributedMAP.compute().
 *
 * Peer blocks.  yh_peer_alloc is the one allocating call of this library (a block other
 * processes map must come from cudaMalloc): it returns a zeroed device block and its
 * 64-byte CUDA IPC handle; the other ranks of the box map it with yh_peer_open.  A rank's
 * block is yh_peer_block_bytes(capacity) long: control words, then capacity keys at
 * yh_peer_key_offset and capacity flag words at yh_peer_flags_offset.
 *
 * yh_exchange_records (collective over `world` ranks, every rank with the same capacity
 * and the same `epoch`: either > 0 and growing by one per call, or 0 on every rank = "use the
 * call counter the library keeps in the rank's own block" -- the form a CUDA-graph capture
 * needs, where kernel arguments are frozen between replays): peer_block_host[r] is rank
 * r's block as mapped in THIS process (the rank's own allocation at [rank]).  On return
 * (stream order) this rank's block holds, in rank order and inside a rank in input
 * order, every record whose class this rank owns; info[0] = their count (use it as M_dev
 * of yh_ap_reduce with M = capacity), info[1] = images summed over ranks, info[2] = 1 if a
 * receive buffer would have overflowed (nothing was moved; identical on every rank --
 * fall back to the NCCL exchange and re-create larger blocks), info[3] = 1 if a wait
 * on a peer gave up after 4 s.  gt_count_global[4*C] = the ranks' gt_count_local summed.
 * Between two calls the ranks must be ordered by some collective that completes after the
 * receive buffer was consumed (yh_peer_sum of the AP table is one). */
/* yh_exchange_images (collective, same blocks / epoch / info / gt_count conventions as
 * yh_exchange_records): compaction, partition by class owner and the all-to-all in ONE pass over
 * the fixed-stride records of yh_map_update -- n_chunks batches in update() order, batch c holding
 * n_images[c] images with n_keep[c][i] records at rec_key[c][i*K ..] / rec_flags[c][i*K ..]
 * (host arrays of device pointers).  The records of an image are sorted by class, so they already
 * are one contiguous run per owner: the call counts those runs (8 counters per image), scans them
 * and copies every run once, straight into its owner's receive buffer.  Replaces
 * yh_compact_records + yh_exchange_records (three reads and writes of every record) for
 * world <= 8; the receive buffers end up with the same contents in the same order.  No reference
 * counterpart (single-process). */
size_t yh_exchange_images_workspace_bytes(int64_t n_images_total, int n_chunks);
int yh_exchange_images(int n_chunks, const int32_t *const *n_keep, const uint64_t *const *rec_key,
                       const uint64_t *const *rec_flags, const int *n_images, int K,
                       const uint8_t *owner_of_class_host, int C, int world, int rank,
                       void *const *peer_block_host, int64_t capacity, unsigned epoch,
                       const int32_t *gt_count_local, int32_t *gt_count_global, int64_t *info,
                       void *workspace, size_t workspace_bytes, void *stream);

/* yh_peer_sum (collective, same epoch rule): vec[n] (device doubles, n <= 4*32*13 + 4) becomes
 * the sum over the ranks, in rank order, through the peer blocks -- the all-reduce of the
 * packed AP table without a second library.  vec[err_index] is set to 1 if a peer did not
 * show up.  A rank past this call knows every peer has consumed its receive buffer, which
 * is the ordering yh_exchange_records asks for between two calls. */
int yh_peer_sum(double *vec, int n, int err_index, int world, int rank, void *const *peer_block_host,
                unsigned epoch, void *stream);
int yh_peer_alloc(size_t bytes, void **ptr, void *handle64);
int yh_peer_open(const void *handle64, void **ptr);
/* Waits on a peer's flag give up after wait_ms milliseconds of device time (default 4000; a rank's own block). */
int yh_peer_set_wait_ms(void *own_block, unsigned wait_ms);
int yh_peer_close(void *ptr);
int yh_peer_free(void *ptr);
size_t yh_peer_block_bytes(int64_t capacity);
size_t yh_peer_key_offset(int64_t capacity);
size_t yh_peer_flags_offset(int64_t capacity);
size_t yh_exchange_workspace_bytes(int64_t M);
int yh_exchange_records(const uint64_t *key, const uint64_t *flags, int64_t M, const int64_t *M_dev,
                        const uint8_t *owner_of_class_host, int C, int world, int rank,
                        void *const *peer_block_host, int64_t capacity, unsigned epoch,
                        const int32_t *gt_count_local, int64_t n_images_local, int32_t *gt_count_global,
                        int64_t *info, void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------------ target encoder ----
 * Replaces VOCDetectionYOLO._encode_target (src/yolo/dataset.py:487-532) for a batch:
 * image n owns boxes [offsets[n], offsets[n+1]) of `boxes` (fp64 x_center, y_center, w, h,
 * normalised) and `cls`; out[n] is its (S, S, B*5 + C) fp32 target: the first box of a
 * cell in list order fills slot 0 (cell-relative x, y; w; h; confidence 1) and the class
 * one-hot, later boxes of the same cell are dropped.  bad[n] (nullable) = 1 where the
 * reference would raise (non-finite centre, cell or class index out of range). */
int yh_encode_targets(const double *boxes, const int32_t *cls, const int64_t *offsets, int N, int S,
                      int B, int C, float *out, uint8_t *bad, void *stream);

/* ------------------------------------------------ image preprocessing ----
 * Replaces `self.transform(image)` of YOLOInference (src/yolo/inference.py:57-66, 88-101):
 * torchvision Resize((out_h, out_w)) on a PIL RGB image (Pillow's bilinear resample: 22-bit
 * fixed-point taps, horizontal then vertical pass, 8-bit rounding in between), ToTensor
 * (/ 255) and Normalize((x - mean) / std), all in float32.  rgb = (N, H, W, 3) uint8 on the
 * device (N images of one size, 4-byte aligned), out = (N, 3, out_h, out_w) float32,
 * bit-identical to the reference's tensor for every image; three launches for the batch.
 * mean3 / std3 are HOST arrays of three floats.  Uses `workspace` (256-byte aligned) of
 * yh_preprocess_workspace_bytes(N, H, W, out_h, out_w); 0 = shape not supported. */
size_t yh_preprocess_workspace_bytes(int N, int H, int W, int out_h, int out_w);
int yh_preprocess_rgb8(const uint8_t *rgb, int N, int H, int W, int out_h, int out_w, const float *mean3,
                       const float *std3, float *out, void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------- list-form entry points ----
 * The reference's helpers also accept arbitrary Python lists (tests call
 * _apply_nms / non_max_suppression / _calculate_iou / iou on hand-built boxes,
 * and compute() runs on whatever all_predictions holds).  These take CSR lists
 * of fp64 boxes (x, y, w, h): list i owns [off[i], off[i+1]). */
int yh_iou_pairs(const double *box_a, const double *box_b, int64_t n, int iou_mode, double *out,
                 void *stream);
/* keep_idx[off[i] + q] = index (within list i) of the q-th kept box; n_keep[i].
 * max_len = longest list (host-known); at most 1024 boxes per list in this build. */
int yh_nms_lists(const int64_t *off, int n_lists, const int32_t *cls, const double *conf,
                 const double *box, double nms_thr, int iou_mode, int order, int32_t *n_keep,
                 int32_t *keep_idx, int max_len, void *stream);
/* TP flags for list-form detections / ground truths of the same n_lists images.
 * Detections may be in any per-image order: "earlier" is decided as the global
 * stable sort does (higher confidence, or equal confidence and earlier in the
 * list).  Emits the same rec_key / rec_flags as yh_map_update (dense, CSR order)
 * and adds to gt_count[4*C].  max_dets = most detections in one image (<= 1024). */
int yh_match_lists(const int64_t *det_off, const int32_t *det_cls, const double *det_conf,
                   const double *det_box, const int64_t *gt_off, const int32_t *gt_cls,
                   const double *gt_box, int n_lists, int C, const double *thresholds_host, int T,
                   uint64_t *rec_key, uint64_t *rec_flags, int32_t *gt_count, int max_dets,
                   void *stream);

/* ---------------------------------------------------------------------------------------------------------------
 * The head's last layer fused with the loss.  Replaces, for S=7, B=2, C=20, in_features=4096 and
 * fp16 / bf16 operands (what autocast makes of the layer: src/yolo/training/trainer.py:70, 165):
 *     x = nn.Linear(4096, S*S*(B*5+C))(x); x = x.view(-1, S, S, B*5+C)      src/yolo/models.py:326-331, 345-348
 *     loss, loss_dict = YOLOLoss(...)(x, targets); d loss / d x               src/yolo/loss.py:55-172 + autograd
 * by ONE tcgen05 / TMEM / TMA kernel whose epilogue evaluates the loss on the accumulator tile: the [M, 1470]
 * prediction tensor is not written to memory unless pred_out asks for it.
 *   x [M, 4096], weight [1470, 4096] (nn.Linear's own layout), both `dtype`, 16-byte aligned: YH_F16 / YH_BF16, or fp32
 *   operands for the reference's default (non-AMP) step -- YH_F32 = 3xTF32 (each operand split into the 19 bits TF32
 *   keeps and the rest, three tensor-core passes: fp32's accuracy), YH_TF32 = one TF32 pass (what torch does under
 *   allow_tf32); fp32 operands take an fp32 gradient and ignore round_pred;
 *   bias [1470] fp32 or NULL; tgt [M, 7, 7, 30] fp32, or NULL = plain Linear (pred_out required, parts / grad unused);
 *   round_pred != 0: the prediction is rounded to `dtype` before the loss sees it, as the autocast Linear's output is;
 *   parts[5] fp32 = total, coord, conf_obj, conf_noobj, class (loss.py:164-170);
 *   grad [M, 1470] = grad_scale * d total / d prediction in `grad_dtype` (= dtype, or YH_F32), or NULL -- grad_scale is
 *   the GradScaler factor: fp16 cannot hold 2 d / M for a large M;
 *   pred_out [M, 1470] fp32 or NULL (feeds yh_map_update / yh_decode_nms directly).
 *   workspace: yh_head_linear_loss_workspace_bytes() = a ticket word, one partial sum set per SM and 24 MB in which a YH_F32 call
 *   keeps the part of the weight that TF32 drops (split off once per call); needed when tgt != NULL or dtype == YH_F32.
 * The caller finishes the layer's backward with two library GEMMs (d x = grad @ weight, d weight = grad^T @ x).
 * Accumulation is fp32; the loss arithmetic is yh_loss_fwd_bwd's on the same predictions, bit for bit; the five sums are
 * folded in a fixed order for a given GPU (one partial per SM).  Other shapes / dtypes: YH_ERR_UNSUPPORTED. */
size_t yh_head_linear_loss_workspace_bytes(void);
int yh_head_linear_loss(const void *x, const void *weight, const float *bias, int dtype, const float *tgt, int M,
                        int S, int B, int C, int in_features, float lambda_coord, float lambda_noobj,
                        float grad_scale, int round_pred, float *parts, void *grad, int grad_dtype,
                        float *pred_out, void *workspace, size_t workspace_bytes, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* YOLOHEAD_H */
