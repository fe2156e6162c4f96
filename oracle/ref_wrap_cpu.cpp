// Builds the reference's own CPU BEV IoU (compiled from where it lies, unmodified) behind a C ABI.
// REF_CPU_SRC is passed by oracle/build_oracle.py as the absolute path of
// mmdet3d/ops/pcdet_nms/src/iou3d_cpu.cpp.
#include REF_CPU_SRC
extern "C" void ref_iou_bev_matrix_cpu(const float* a, int na, const float* b, int nb, float* out) {
  for (int i = 0; i < na; i++)
    for (int j = 0; j < nb; j++) out[(long)i * nb + j] = iou_bev(a + i * 7, b + j * 7);
}
