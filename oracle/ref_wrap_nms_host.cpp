// Builds the reference's own pcdet_nms HOST code (mmdet3d/ops/pcdet_nms/src/iou3d_nms.cpp, compiled from where it lies,
// unmodified, against oracle/ref_stubs_host) behind a C ABI, so that bench.py can time the reference's complete NMS call -
// cudaMalloc + mask kernel + blocking D2H copy + host greedy loop + cudaFree (iou3d_nms.cpp:90-136 / :139-186) - on the
// same GPU as the product's batched kernel.  REF_NMS_SRC = absolute path of that file.
#include REF_NMS_SRC
extern "C" int ref_nms_gpu_call(const float* boxes_dev, int boxes_num, long* keep_host, float thr, int rotated) {
  at::Tensor boxes{const_cast<float*>(boxes_dev), boxes_num, 7, true};
  at::Tensor keep{keep_host, boxes_num, 1, false};
  return rotated ? nms_gpu(boxes, keep, thr) : nms_normal_gpu(boxes, keep, thr);
}
