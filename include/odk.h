/* odk.h — C-ABI of libodk.so: the B200 (sm_100a) detection hot path.
 *
 * The reference (ak110/object_detector, pytoolkit @ d542569) has no FFI: its boundary for this
 * path is a set of Python call signatures.  Each entry point below names the reference function
 * it replaces (paths relative to pytoolkit/pytoolkit/).  The Python mirror in
 * object_detector_b200/ binds these with ctypes; INTEGRATION.md shows the stub a pytoolkit
 * maintainer would add.
 *
 * Conventions
 *  - every pointer marked [dev] is a DEVICE pointer owned by the caller (a torch tensor's
 *    data_ptr()); [host] pointers are host memory, read during the call only;
 *  - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).  Launches are
 *    asynchronous: no entry point synchronises the stream;
 *  - every function returns 0 on success or a negative odk_status; the message is retrievable
 *    per host thread with odk_last_error().  Safe to call from many host threads on distinct
 *    streams (generator.py:160-166 calls encode_truth from a thread pool).  State the library
 *    keeps besides the anchors handles, all of it per host thread or diagnostic:
 *      * odk_assign_encode / odk_assign_encode_compact fork part of their work onto LIBRARY-OWNED
 *        side streams (a small per-host-thread, per-device pool of non-blocking streams + events,
 *        created on first use, never destroyed) and join them back with events before returning
 *        to the caller's stream order; the fork/join is capturable into a CUDA graph.  Work is
 *        therefore ordered with respect to `stream`, but not every kernel runs ON `stream`;
 *      * odk_profile_enable() switches on per-host-thread CUDA-event timing of every launch; while it is
 *        on, odk_assign_encode runs its two chains one after the other on `stream` (no fork), so that
 *        every event pair brackets one kernel running alone;
 *      * odk_timeline_set() installs a PROCESS-WIDE device pointer (one per device context) that
 *        instrumented kernels write %globaltimer stamps to — a diagnostic for overlap studies,
 *        never set on the product path;
 *      * kernel tuning knobs are read once from ODK_* environment variables (diagnostics);
 *  - the library allocates device memory only inside odk_anchors_create (freed by
 *    odk_anchors_destroy).  Scratch memory is passed in by the caller (odk_*_workspace_bytes).
 */
#ifndef ODK_H_
#define ODK_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  ODK_OK = 0,
  ODK_ERR_INVALID_ARGUMENT = -1, /* shape / range / alignment: the reference's assert on shapes       */
  ODK_ERR_CUDA = -2,             /* a CUDA runtime call failed (message has cudaGetErrorString)       */
  ODK_ERR_WORKSPACE = -3,        /* workspace smaller than odk_*_workspace_bytes()                    */
  ODK_ERR_UNSUPPORTED = -4       /* size outside what the kernels are built for                       */
} odk_status;

/* Per-image status words written by odk_assign / odk_assign_encode into `status[B]` [dev]
 * (they replace the reference's AssertionErrors, which are raised per image):            */
#define ODK_IMG_OK 0
#define ODK_IMG_ENCODE_ERROR 1   /* 'Encode error: {centre}'  od_pb.py:332,338                       */
#define ODK_IMG_BAD_CLASS 2      /* assert in_range(classes, 0, C)  od_pb.py:304                      */
#define ODK_IMG_BAD_COUNT 3      /* gt_count[b] outside [0, Gmax]                                     */

typedef struct odk_anchors odk_anchors; /* opaque: device copies of the PriorBoxes tables */

const char* odk_version(void);
const char* odk_last_error(void);

/* PriorBoxes._create_pb tables (dl/od_pb.py:125-184), built on the host by the Python mirror and
 * uploaded once.  Anchor order is the reference's: maps in `map_sizes` order, then pattern, then
 * row-major y*m+x (SURVEY A-1); n must equal n_patterns * sum(m*m).  pb_sizes stays fp64 because
 * encode_locs divides in fp64 (od_pb.py:212-214); the fp32 decode scale float32(0.10*pb_sizes)
 * (od_pb.py:221 as evaluated in the fp32 graph) is derived here. */
int odk_anchors_create(odk_anchors** out, int n, int num_classes, int n_maps,
                       const int32_t* map_sizes /*[host n_maps]*/, int n_patterns,
                       const float* pb_locs /*[host n,4]*/, const float* pb_grid /*[host n,4]*/,
                       const double* pb_sizes /*[host n,2]*/, const uint8_t* pb_mask /*[host n]*/);
int odk_anchors_destroy(odk_anchors* a);
int odk_anchors_count(const odk_anchors* a);
int odk_anchors_num_classes(const odk_anchors* a);
/* Statistics of the valid-anchor span table the fused predict path reads through (anchors with pb_mask = 0 are
 * never fetched there: conf = x * pb_mask = 0 can not pass conf_thr >= 0, od_net.py:183-184):
 * out4 = {tile height, spans in the table, anchor rows covered by the spans, anchors with pb_mask = 1}. */
int odk_anchors_span_stats(const odk_anchors* a, int32_t* out4 /*[host 4]*/);

/* ml.compute_iou (ml.py:502-519) on fp32 boxes: out[i*nb+j] = IoU(a[i], b[j]). */
int odk_compute_iou(const float* a /*[dev na,4]*/, int na, const float* b /*[dev nb,4]*/, int nb,
                    float* out /*[dev na,nb]*/, void* stream);

/* PriorBoxes.encode_locs (od_pb.py:212-214): out[k] = (boxes[k] - pb_locs[ix[k]]) / sizes / 0.10,
 * fp64 result like the reference. */
int odk_encode_locs(const odk_anchors* a, int k, const float* boxes /*[dev k,4]*/,
                    const int32_t* ix /*[dev k]*/, double* out /*[dev k,4]*/, void* stream);

/* PriorBoxes.decode_locs as the fp32 Keras Lambda (od_pb.py:216-222, od_net.py:189):
 * out = clip(pred * float32(0.10*sizes) + pb_locs, 0, 1) over `rows` = B*N anchors rows. */
int odk_decode_locs(const odk_anchors* a, int64_t rows, const float* pred /*[dev rows,4]*/,
                    float* out /*[dev rows,4]*/, void* stream);

/* PriorBoxes.decode_locs with xp=np (od_pb.py:216-222 as NumPy evaluates it; the call check_assign.py:27 makes):
 * pb_sizes is fp64 there, so out = clip(pred * (0.10 * tile(pb_sizes,2)) + pb_locs, 0, 1) entirely in fp64. */
int odk_decode_locs_f64(const odk_anchors* a, int64_t rows, const double* pred /*[dev rows,4]*/,
                        double* out /*[dev rows,4]*/, void* stream);

/* ChannelArgMax / ChannelMax as stand-alone layers (layers.py:69-98): per row of `channels` values the index of the
 * first maximum (int64, like K.argmax) and/or the maximum.  Either output may be NULL.  (odk_decode_score and
 * odk_decode_nms fuse the same scan; these tensors never exist on the hot path.) */
int odk_channel_argmax_max(int64_t rows, int channels, const float* x /*[dev rows,channels]*/,
                           int64_t* arg /*[dev rows] or NULL*/, float* mx /*[dev rows] or NULL*/, void* stream);

/* PriorBoxes._assign_boxes (od_pb.py:315-363) for a padded batch.  Outputs per image b with
 * G = gt_count[b]: pb_indices[b, 0:G] ascending anchor ids, gt_indices[b, 0:G] the GT each one
 * owns (both -1 beyond G), status[b] one of ODK_IMG_*. */
size_t odk_assign_workspace_bytes(const odk_anchors* a, int batch, int gmax); /* scratch for odk_assign / odk_assign_encode */
int odk_assign(const odk_anchors* a, int batch, const float* gt_boxes /*[dev B,Gmax,4]*/,
               const int32_t* gt_count /*[dev B]*/, int gmax,
               int64_t* pb_indices /*[dev B,Gmax]*/, int64_t* gt_indices /*[dev B,Gmax]*/,
               int32_t* status /*[dev B]*/, void* workspace /*[dev]*/, size_t workspace_bytes,
               void* stream);

/* PriorBoxes.encode_truth (od_pb.py:299-313) = odk_assign + the y_true writer.
 * y_true[b, n, :] = [mask, weight, obj, onehot(C), locs(4)] fp32, written in full (no pre-zeroing
 * needed).  pb_valids (optional, may be NULL) receives _assign_boxes' third result. */
int odk_assign_encode(const odk_anchors* a, int batch, const float* gt_boxes /*[dev B,Gmax,4]*/,
                      const int32_t* gt_classes /*[dev B,Gmax]*/,
                      const int32_t* gt_count /*[dev B]*/, int gmax,
                      float* y_true /*[dev B,N,3+C+4]*/, int64_t* pb_indices /*[dev B,Gmax]*/,
                      int64_t* gt_indices /*[dev B,Gmax]*/, uint8_t* pb_valids /*[dev B,N] or NULL*/,
                      int32_t* status /*[dev B]*/, void* workspace /*[dev]*/, size_t workspace_bytes,
                      void* stream);

/* PriorBoxes.encode_truth (od_pb.py:299-313) with a COMPACT result, for callers whose consumer sits on the host side
 * of PCIe (the reference's generator hands y_true to Keras through host memory, generator.py:227-240): the dense
 * (N, 3+C+4) tensor is > 99.7 % zeros plus one repeated constant, so only what defines it leaves the kernel:
 *   valid_bits[b, w]  bit a of word a>>5 = pb_valids[a] (third result of _assign_boxes, od_pb.py:355-361); bits >= N are 0
 *   pos_rows[b, k, :] the y_true row of anchor pb_indices[b, k] for k < G (zeros beyond G): [w0, 1/G, 1, onehot, locs]
 * With w0 = float32(1/max(1,G)):  y_true[b] = zeros; y_true[b, a, 0] = w0 where valid_bits says so;
 * y_true[b, pb_indices[b, k], :] = pos_rows[b, k, :]   — bit-identical to odk_assign_encode's output
 * (object_detector_b200.od_pb.expand_truth does exactly this).  Failed images: all bits and rows zero.
 * Same workspace as odk_assign_encode. */
int odk_assign_encode_compact(const odk_anchors* a, int batch, const float* gt_boxes /*[dev B,Gmax,4]*/,
                              const int32_t* gt_classes /*[dev B,Gmax]*/, const int32_t* gt_count /*[dev B]*/,
                              int gmax, float* pos_rows /*[dev B,Gmax,3+C+4]*/,
                              uint32_t* valid_bits /*[dev B,ceil(N/32)]*/, int64_t* pb_indices /*[dev B,Gmax]*/,
                              int64_t* gt_indices /*[dev B,Gmax]*/, int32_t* status /*[dev B]*/,
                              void* workspace /*[dev]*/, size_t workspace_bytes, void* stream);

/* PriorBoxes.loss / loss_obj / loss_clf / loss_loc and the rec_bg / rec_obj / acc_clf metrics
 * (od_pb.py:365-431; binary focal loss losses.py:22-30, Keras categorical_crossentropy and
 * mean_squared_error) — forward values per image.  y_true, y_pred: [dev B,N,3+C+4] fp32 in the
 * encode_truth layout.  out[b, :] = {loss, loss_obj, loss_clf, loss_loc, rec_bg, rec_obj, acc_clf,
 * number of positive anchors} as fp64 (per-row arithmetic fp32 like the TF graph, sums fp64).
 * PARITY UNPINNED (TensorFlow absent from the reference tree); 1e-5 relative against the oracle. */
size_t odk_loss_workspace_bytes(const odk_anchors* a, int batch);
int odk_loss(const odk_anchors* a, int batch, const float* y_true /*[dev B,N,3+C+4]*/,
             const float* y_pred /*[dev B,N,3+C+4]*/, double* out /*[dev B,8]*/,
             void* workspace /*[dev]*/, size_t workspace_bytes, void* stream);

/* Gradient of PriorBoxes.loss (od_pb.py:365-373: loss_obj + loss_clf + loss_loc) with respect to y_pred — what
 * TensorFlow's autodiff derives from the reference's graph ops (tf.clip_by_value passes gradients only inside the
 * clip range).  grad[b, n, :] = upstream[b] * dloss[b]/dy_pred[b, n, :]; upstream may be NULL (= 1). */
int odk_loss_grad(const odk_anchors* a, int batch, const float* y_true /*[dev B,N,3+C+4]*/,
                  const float* y_pred /*[dev B,N,3+C+4]*/, const float* upstream /*[dev B] or NULL*/,
                  float* grad /*[dev B,N,3+C+4]*/, void* stream);

/* encode_truth (od_pb.py:299-313) followed by loss (od_pb.py:365-431) with y_true never written:
 * odk_assign, then a pass that reads one float per negative anchor row of y_pred and whole rows
 * only for the positives.  Same outputs as odk_loss; pb_indices / gt_indices / status as odk_assign.
 * Images whose status is not ODK_IMG_OK contribute an all-zero y_true (as odk_assign_encode). */
size_t odk_assign_loss_workspace_bytes(const odk_anchors* a, int batch, int gmax);
int odk_assign_loss(const odk_anchors* a, int batch, const float* gt_boxes /*[dev B,Gmax,4]*/,
                    const int32_t* gt_classes /*[dev B,Gmax]*/, const int32_t* gt_count /*[dev B]*/,
                    int gmax, const float* y_pred /*[dev B,N,3+C+4]*/, double* out /*[dev B,8]*/,
                    int64_t* pb_indices /*[dev B,Gmax]*/, int64_t* gt_indices /*[dev B,Gmax]*/,
                    int32_t* status /*[dev B]*/, void* workspace /*[dev, 256-byte aligned]*/,
                    size_t workspace_bytes, void* stream);

/* odk_assign_loss plus the gradient of the loss w.r.t. y_pred in the same pass (what a training step needs from
 * encode_truth + loss + autodiff), still without y_true: grad[b, n, :] = upstream[b] * dloss[b]/dy_pred[b, n, :],
 * all zero except the objectness column and the positive rows; upstream may be NULL (= 1). */
int odk_assign_loss_grad(const odk_anchors* a, int batch, const float* gt_boxes /*[dev B,Gmax,4]*/,
                         const int32_t* gt_classes /*[dev B,Gmax]*/, const int32_t* gt_count /*[dev B]*/,
                         int gmax, const float* y_pred /*[dev B,N,3+C+4]*/, const float* upstream /*[dev B] or NULL*/,
                         double* out /*[dev B,8]*/, float* grad /*[dev B,N,3+C+4]*/,
                         int64_t* pb_indices /*[dev B,Gmax]*/, int64_t* gt_indices /*[dev B,Gmax]*/,
                         int32_t* status /*[dev B]*/, void* workspace /*[dev, 256-byte aligned]*/,
                         size_t workspace_bytes, void* stream);

/* ChannelArgMax + ChannelMax + _objconf + decode_locs (layers.py:69-98, od_net.py:177-189):
 * classes = argmax_c clfs, confs = objs*max_c clfs*pb_mask, boxes = decode(locs). */
int odk_decode_score(const odk_anchors* a, int batch, const float* objs /*[dev B,N,1]*/,
                     const float* clfs /*[dev B,N,C]*/, const float* locs /*[dev B,N,4]*/,
                     int32_t* classes /*[dev B,N]*/, float* confs /*[dev B,N]*/,
                     float* boxes /*[dev B,N,4]*/, void* stream);

/* Scratch size for odk_nms_layer / odk_decode_nms / odk_nms. */
size_t odk_nms_workspace_bytes(int batch, int n, int num_classes, int top_k);

/* layers.nms() -> NMS._process_image (layers.py:925-970), nms_all_threshold=None branch:
 * per class: candidates classes==c & confs > conf_thr, greedy NMS, then top_k over all classes,
 * zero padding.  out[b, r, :] = [class, conf, x1, y1, x2, y2]; kept_index[b, r] = anchor id or -1.
 * iou_ge = 1: the reference's NumPy NMS semantics (ml.py:589-625: suppress if IoU >= thr, boxes
 * with w<=0 or h<=0 skipped, length check quirk) — the bit-exact, pinned mode.
 * iou_ge = 0: TensorFlow-1.10-style (suppress if IoU > thr, no zero-area skip, exact cut) —
 * PARITY UNPINNED (TF source absent from the reference tree). */
int odk_nms_layer(int batch, int n, int num_classes, const int32_t* classes /*[dev B,N]*/,
                  const float* confs /*[dev B,N]*/, const float* boxes /*[dev B,N,4]*/,
                  float conf_thr, float iou_thr, int top_k, int iou_ge,
                  float* out /*[dev B,top_k,6]*/, int32_t* kept_index /*[dev B,top_k]*/,
                  void* workspace /*[dev]*/, size_t workspace_bytes, void* stream);

/* Fused odk_decode_score + odk_nms_layer (od_net.py:186-190 in one pass): the dense
 * classes/confs/boxes intermediates never reach HBM. */
int odk_decode_nms(const odk_anchors* a, int batch, const float* objs, const float* clfs,
                   const float* locs, float conf_thr, float iou_thr, int top_k, int iou_ge,
                   float* out /*[dev B,top_k,6]*/, int32_t* kept_index /*[dev B,top_k]*/,
                   void* workspace /*[dev]*/, size_t workspace_bytes, void* stream);

/* strict_nms=True (od_net.py:175 -> layers.py:890 nms_all_threshold=0.5): the per-class limit doubles
 * to 2*top_k (layers.py:940) and a second, class-agnostic greedy NMS at nms_all_thr runs over the
 * class-concatenated survivors in descending score (layers.py:953-957), cut at top_k.  Both stages
 * run inside the same per-image sweep.  Equal scores are visited lower class first in the second
 * stage (tie policy of this build; identical to the reference for distinct scores).  Needs the
 * larger workspace below. */
size_t odk_nms_strict_workspace_bytes(int batch, int n, int num_classes, int top_k);
int odk_nms_layer_strict(int batch, int n, int num_classes, const int32_t* classes /*[dev B,N]*/,
                         const float* confs /*[dev B,N]*/, const float* boxes /*[dev B,N,4]*/,
                         float conf_thr, float iou_thr, float nms_all_thr, int top_k, int iou_ge,
                         float* out /*[dev B,top_k,6]*/, int32_t* kept_index /*[dev B,top_k]*/,
                         void* workspace /*[dev]*/, size_t workspace_bytes, void* stream);
int odk_decode_nms_strict(const odk_anchors* a, int batch, const float* objs, const float* clfs,
                          const float* locs, float conf_thr, float iou_thr, float nms_all_thr,
                          int top_k, int iou_ge, float* out /*[dev B,top_k,6]*/,
                          int32_t* kept_index /*[dev B,top_k]*/, void* workspace /*[dev]*/,
                          size_t workspace_bytes, void* stream);

/* ml.non_maximum_suppression (ml.py:589-625) on one fp32 box set: writes the kept input indices
 * in descending-score order to out_index[0:*out_count] (capacity max(top_k,2)). */
int odk_nms(const float* boxes /*[dev n,4]*/, const float* scores /*[dev n]*/, int n, int top_k,
            float iou_thr, int iou_ge, int32_t* out_index /*[dev max(top_k,2)]*/,
            int32_t* out_count /*[dev 1]*/, void* workspace, size_t workspace_bytes, void* stream);

/* odk_nms without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same inputs, outputs and semantics as odk_nms; the scratch is a static device array of
 * 640 KB inside the library (held on every GPU that loads it until the process exits), which limits n to
 * odk_nms_static_max_boxes() = 65 536 boxes: larger n returns ODK_ERR_UNSUPPORTED with a message naming the limit.
 * The call makes no cudaMalloc / cudaFree / allocating library call.  All calls on a device share the one array, which
 * keeps the previous call's contents (nothing in it needs to start at zero): calls that OVERLAP ON THE GPU — different
 * streams without ordering between them — race; order them on one stream or with events. */
size_t odk_nms_static_max_boxes(void);
int odk_nms_static(const float* boxes /*[dev n,4]*/, const float* scores /*[dev n]*/, int n, int top_k,
                   float iou_thr, int iou_ge, int32_t* out_index /*[dev max(top_k,2)]*/,
                   int32_t* out_count /*[dev 1]*/, void* stream);

/* data.voc.evaluate (data/voc.py:165-197 -> chainercv eval_detection_voc; PARITY UNPINNED).
 * dets: the NMS layer output for n_img images; rows with conf < conf_thr are dropped
 * (od.py:265).  gt_offsets[n_img+1] indexes the flat GT arrays.  Boxes are converted to integer
 * pixels with round-half-even of the fp64 product (ml.py:62-65,184-186). */
size_t odk_voc_ap_workspace_bytes(int n_img, int top_k, int num_classes, int gt_max_per_image);
int odk_voc_ap(int n_img, int top_k, int num_classes, const float* dets /*[dev n_img,top_k,6]*/,
               float conf_thr, const int32_t* img_wh /*[dev n_img,2] (width, height)*/,
               const float* gt_boxes /*[dev n_gt,4] normalised*/, const int32_t* gt_classes /*[dev n_gt]*/,
               const uint8_t* gt_difficult /*[dev n_gt]*/, const int32_t* gt_offsets /*[dev n_img+1]*/,
               int gt_max_per_image, double* ap_area /*[dev C]*/, double* ap07 /*[dev C]*/,
               void* workspace, size_t workspace_bytes, void* stream);

/* data.coco.evaluate (data/coco.py:75-99 -> chainercv eval_detection_coco -> pycocotools COCOeval 'bbox';
 * PARITY UNPINNED: neither package is vendored or pinned by the reference).  dets / img_wh / gt_boxes / gt_classes /
 * gt_offsets as in odk_voc_ap; gt_areas = ObjectsAnnotation.areas (the annotation's area, used for the area ranges),
 * gt_crowd = crowdeds.  iou_thrs[10], rec_thrs[101], area_ranges[4][2] are HOST arrays (pycocotools Params: the
 * caller passes np.linspace's values so that threshold comparisons are bit-identical).  Outputs are COCOeval.eval's
 * arrays over ALL class ids 0..C-1: precision[T=10,R=101,C,A=4,M=3], recall[T,C,A,M], -1 where undefined
 * (maxDets = 1, 10, 100; at most 256 GT boxes per image). */
size_t odk_coco_ap_workspace_bytes(int n_img, int top_k, int num_classes);
int odk_coco_ap(int n_img, int top_k, int num_classes, const float* dets /*[dev n_img,top_k,6]*/, float conf_thr,
                const int32_t* img_wh /*[dev n_img,2]*/, const float* gt_boxes /*[dev n_gt,4] normalised*/,
                const int32_t* gt_classes /*[dev n_gt]*/, const double* gt_areas /*[dev n_gt] pixels^2*/,
                const uint8_t* gt_crowd /*[dev n_gt]*/, const int32_t* gt_offsets /*[dev n_img+1]*/,
                int gt_max_per_image, const double* iou_thrs /*[host 10]*/, const double* rec_thrs /*[host 101]*/,
                const double* area_ranges /*[host 4,2]*/, double* precision /*[dev 10,101,C,4,3]*/,
                double* recall /*[dev 10,C,4,3]*/, void* workspace /*[dev, 256-byte aligned]*/,
                size_t workspace_bytes, void* stream);

/* ml.compute_scores (ml.py:352-394) for n_thr confidence thresholds at once (ml.search_conf_threshold,
 * ml.py:328-340, sweeps 50 of them): counts[t, 0/1/2, c] = TP / FP / FN of class c at conf_thr[t].
 * dets / GT arrays as in odk_voc_ap (boxes normalised fp32, IoU in fp32 like ml.compute_iou). */
int odk_match_counts(int n_img, int top_k, int num_classes, const float* dets /*[dev n_img,top_k,6]*/,
                     const float* gt_boxes /*[dev n_gt,4]*/, const int32_t* gt_classes /*[dev n_gt]*/,
                     const uint8_t* gt_difficult /*[dev n_gt]*/, const int32_t* gt_offsets /*[dev n_img+1]*/,
                     int n_thr, const float* conf_thr /*[dev n_thr]*/, float iou_thr,
                     int32_t* counts /*[dev n_thr,3,C]*/, void* stream);

/* ml.od_confusion_matrix (ml.py:397-451): cm[(C+1),(C+1)] int32, rows = ground-truth class (last row:
 * nothing there), columns = predicted class (last column: missed).  dets as for odk_match_counts
 * ([n_img, top_k, 6] rows [class, conf, x1,y1,x2,y2], real rows first, class < 0 = padding).  The
 * reference's positional use of the mask at ml.py:421 is reproduced (see csrc/scores.cu). */
int odk_confusion_matrix(int n_img, int top_k, int num_classes, const float* dets /*[dev n_img,top_k,6]*/,
                         const float* gt_boxes /*[dev n_gt,4]*/, const int32_t* gt_classes /*[dev n_gt]*/,
                         const int32_t* gt_offsets /*[dev n_img+1]*/, float conf_thr, float iou_thr,
                         int32_t* cm /*[dev (C+1)*(C+1)]*/, void* stream);

/* ObjectsPrediction.is_match (ml.py:159-182) per image -> match[n_img] (1 = the detections with
 * conf >= conf_thr pair off one-to-one with the ground truth, IoU >= iou_thr, same classes);
 * ml.od_accuracy (ml.py:343-349) is the mean of it. */
int odk_exact_match(int n_img, int top_k, const float* dets /*[dev n_img,top_k,6]*/,
                    const float* gt_boxes /*[dev n_gt,4]*/, const int32_t* gt_classes /*[dev n_gt]*/,
                    const int32_t* gt_offsets /*[dev n_img+1]*/, float conf_thr, float iou_thr,
                    uint8_t* match /*[dev n_img]*/, void* stream);

/* Bounding-box side of the augmentation upstream of encode_truth, for a padded batch of ground truth:
 * RandomZoom._padding / _crop (image.py:321-388) and the letterbox of a failed zoom (image.py:298-314).
 * kind[b]: 0 leave alone, 1 zoom-out (paste at try_xy inside a try_size canvas), 2 zoom-in (crop window
 * try_xy..try_xy+try_size), 3 letterbox (in_size = resized image, try_size = output, try_xy = origin; always
 * accepted).  Each image walks its `tries` attempts in order, as the reference's retry loop does, and takes the
 * first acceptable one: chosen[b] = its index or -1 (boxes then pass through unchanged).  boxes are fp64 holding
 * either fp32 values straight from ObjectsAnnotation (src_f32 = 1: centre and area are then taken in fp32, as NumPy
 * does) or the fp64 result of an earlier step.  keep[b,g] = 0 for boxes the crop drops (image.py:355-363);
 * wide[b] = 1 once the boxes of image b are fp64 values in the reference. */
int odk_bbox_zoom(int batch, int gmax, int tries, const double* boxes /*[dev B,Gmax,4]*/, int src_f32,
                  const int32_t* gt_count /*[dev B]*/, const int32_t* kind /*[dev B]*/,
                  const int64_t* in_size /*[dev B,2] w,h*/, const int64_t* try_size /*[dev B,tries,2]*/,
                  const int64_t* try_xy /*[dev B,tries,2]*/, int out_w, int out_h, double min_object_px,
                  double* boxes_out /*[dev B,Gmax,4]*/, uint8_t* keep /*[dev B,Gmax]*/,
                  int32_t* chosen /*[dev B]*/, uint8_t* wide /*[dev B]*/, void* stream);

/* RandomFlipLR / RandomFlipTB (image.py:433-465) then ObjectsAnnotation.rot90 (ml.py:102-112), in that order
 * (od.py:160-165), followed by the order-preserving removal of boxes with keep = 0.  ops[b]: bit 0 flip
 * left-right, bit 1 flip top-bottom, bits 2-3 the number of quarter turns.  1-x is taken in fp32 for images whose
 * wide flag is 0 (NULL: all).  Outputs feed odk_assign_encode directly: boxes_out fp32 [B,Gmax,4], classes_out,
 * count_out (rows beyond the count are zero); boxes64_out (optional) keeps the unrounded values. */
int odk_bbox_flip_rot(int batch, int gmax, const double* boxes /*[dev B,Gmax,4]*/,
                      const uint8_t* keep /*[dev B,Gmax] or NULL*/, const int32_t* gt_classes /*[dev B,Gmax]*/,
                      const int32_t* gt_count /*[dev B]*/, const uint8_t* wide /*[dev B] or NULL*/,
                      const int32_t* ops /*[dev B]*/, double* boxes64_out /*[dev B,Gmax,4] or NULL*/,
                      float* boxes_out /*[dev B,Gmax,4]*/, int32_t* classes_out /*[dev B,Gmax]*/,
                      int32_t* count_out /*[dev B]*/, void* stream);

/* Diagnostics: in-kernel timeline of the seven kernels of a step (candidates, greedy, rows, patch, emit, nms, ignore).
 * buf = 14 device words, word 2*id = earliest CTA entry (initialise to UINT64_MAX), word 2*id+1 = latest CTA exit
 * (initialise to 0), in %globaltimer nanoseconds; NULL switches it off.  The pointer is handed to the kernels as a
 * launch argument (process-wide host state, diagnostics only): a CUDA graph traces what was installed when it was
 * CAPTURED, so install the buffer before capturing and reset its words between replays. */
int odk_timeline_set(unsigned long long* buf /*[dev 14] or NULL*/);

/* Measurement aid (bench.py): when enabled for the calling host thread every kernel launch of the
 * library is bracketed by CUDA events on its stream; odk_profile_read synchronises them and returns,
 * per kernel name, the summed device time and the launch count, then clears the log. */
int odk_profile_enable(int on);
int odk_profile_read(char* names /*[host cap*name_stride]*/, int name_stride, float* total_ms /*[host cap]*/,
                     int32_t* launches /*[host cap]*/, int cap, int32_t* n_out /*[host 1]*/);

#ifdef __cplusplus
}
#endif
#endif /* ODK_H_ */
