// TEST INFRASTRUCTURE (checker, never the product, never timed as the GPU arm).
// pybind11 binding of the REFERENCE's own modulated deformable convolution entry points, compiled
// together with the reference's two unmodified CUDA sources where they lie under /root/reference
// (recipe: oracle/build_ref.py; output: oracle/_ref/, git-ignored, ships to the GPU box).
// Binds exactly the two dispatch functions the reference's vision.cpp:88-95 registers:
//   modulated_deform_conv_forward   deform_conv.h:262-312  -> deform_conv_cuda.cu:804-927
//   modulated_deform_conv_backward  deform_conv.h:314-374  -> deform_conv_cuda.cu:929-1129
#include <torch/extension.h>
#include "deform_conv.h"   // found with -I<reference>/detectron2/layers/csrc/deformable

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("modulated_deform_conv_forward", &detectron2::modulated_deform_conv_forward);
  m.def("modulated_deform_conv_backward", &detectron2::modulated_deform_conv_backward);
}
