// TEST INFRASTRUCTURE (oracle/): python binding for the *reference's* quantizer launchers.
//
// The reference's own csrc/fused/fused.h has six of its eight prototypes commented out
// (/root/reference/csrc/fused/fused.h:19-75) and setup.py never builds the CUDA `_fused`
// extension (/root/reference/setup.py:246-260).  This file only *declares* the launchers that
// are defined in /root/reference/csrc/fused/fused.cu (lines 429, 512, 594, 685, 770, 850, 925,
// 1002) and exposes them to Python, so the unmodified reference source can serve as the GPU
// oracle for bit-exact quantizer parity.  No reference code is copied here.
#include <torch/extension.h>

void quant_per_block_int8_cuda(torch::Tensor input, torch::Tensor output, torch::Tensor scale,
                               float sm_scale, int block_size, int tensor_layout);
void quant_per_block_int8_cuda(torch::Tensor input, torch::Tensor output, torch::Tensor scale,
                               int block_size, int tensor_layout);
void quant_per_block_int8_fuse_sub_mean_cuda(torch::Tensor input, torch::Tensor mean,
                                             torch::Tensor output, torch::Tensor scale,
                                             int block_size, int tensor_layout);
void quant_per_warp_int8_cuda(torch::Tensor input, torch::Tensor output, torch::Tensor scale,
                              int block_size, int warp_block_size, int tensor_layout);
void sub_mean_cuda(torch::Tensor input, torch::Tensor mean, torch::Tensor output, int tensor_layout);
void transpose_pad_permute_cuda(torch::Tensor input, torch::Tensor output, int tensor_layout);
void scale_fuse_quant_cuda(torch::Tensor input, torch::Tensor output, torch::Tensor scale,
                           int num_tokens, float scale_max, int tensor_layout);
void mean_scale_fuse_quant_cuda(torch::Tensor input, torch::Tensor output, torch::Tensor mean,
                                torch::Tensor scale, int num_tokens, float scale_max,
                                int tensor_layout);

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("quant_per_block_int8_sm_scale_cuda",
        py::overload_cast<torch::Tensor, torch::Tensor, torch::Tensor, float, int, int>(
            &quant_per_block_int8_cuda));
  m.def("quant_per_block_int8_cuda",
        py::overload_cast<torch::Tensor, torch::Tensor, torch::Tensor, int, int>(
            &quant_per_block_int8_cuda));
  m.def("quant_per_block_int8_fuse_sub_mean_cuda", &quant_per_block_int8_fuse_sub_mean_cuda);
  m.def("quant_per_warp_int8_cuda", &quant_per_warp_int8_cuda);
  m.def("sub_mean_cuda", &sub_mean_cuda);
  m.def("transpose_pad_permute_cuda", &transpose_pad_permute_cuda);
  m.def("scale_fuse_quant_cuda", &scale_fuse_quant_cuda);
  m.def("mean_scale_fuse_quant_cuda", &mean_scale_fuse_quant_cuda);
}
