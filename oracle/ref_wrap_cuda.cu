// Builds the reference's own CUDA NMS / BEV IoU kernels (compiled from where they lie, unmodified, for sm_100a)
// behind a C ABI.  REF_CUDA_SRC = absolute path of mmdet3d/ops/pcdet_nms/src/iou3d_nms_kernel.cu.
#include REF_CUDA_SRC
extern "C" {
// device pointers; mask: boxes_num * DIVUP(boxes_num, 64) words
int ref_nms_mask(const float* boxes, unsigned long long* mask, int boxes_num, float thr, int rotated) {
  if (rotated) nmsLauncher(boxes, mask, boxes_num, thr); else nmsNormalLauncher(boxes, mask, boxes_num, thr);
  return (int)cudaDeviceSynchronize();
}
int ref_boxes_iou_bev(int na, const float* a, int nb, const float* b, float* out) {
  boxesioubevLauncher(na, a, nb, b, out);
  return (int)cudaDeviceSynchronize();
}
int ref_boxes_overlap_bev(int na, const float* a, int nb, const float* b, float* out) {
  boxesoverlapLauncher(na, a, nb, b, out);
  return (int)cudaDeviceSynchronize();
}
}
