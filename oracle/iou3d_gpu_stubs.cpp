// TEST INFRASTRUCTURE ONLY — link-time stand-ins for the CUDA entry points that the reference's pybind table
// (pcdet/ops/iou3d_nms/src/iou3d_nms_api.cpp) takes the address of.  They are never called: oracle/evolve.py uses only
// boxes_iou_bev_cpu from the reference's own iou3d_cpu.cpp.  Declarations: pcdet/ops/iou3d_nms/src/iou3d_nms.h.
#include <torch/extension.h>
#include <cstdlib>
static int no_gpu_here() { std::abort(); return 0; }
int boxes_aligned_overlap_bev_gpu(at::Tensor, at::Tensor, at::Tensor) { return no_gpu_here(); }
int boxes_overlap_bev_gpu(at::Tensor, at::Tensor, at::Tensor) { return no_gpu_here(); }
int paired_boxes_overlap_bev_gpu(at::Tensor, at::Tensor, at::Tensor) { return no_gpu_here(); }
int boxes_iou_bev_gpu(at::Tensor, at::Tensor, at::Tensor) { return no_gpu_here(); }
int nms_gpu(at::Tensor, at::Tensor, float) { return no_gpu_here(); }
int nms_normal_gpu(at::Tensor, at::Tensor, float) { return no_gpu_here(); }
