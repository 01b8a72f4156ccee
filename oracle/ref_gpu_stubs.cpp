// TEST INFRASTRUCTURE ONLY (oracle/): link-time stubs for the CUDA half of the
// reference's lietorch binding, so that the reference's own CPU sources
// (thirdparty/lietorch/lietorch/src/lietorch.cpp + lietorch_cpu.cpp, compiled where
// they lie under /root/reference) can be built into oracle/_ref/ without nvcc
// (the CUDA half takes ~7 min to compile and is never used by the CPU oracle).
// Signatures follow thirdparty/lietorch/lietorch/include/lietorch_gpu.h:10-46.
#include <torch/extension.h>
#include <vector>

#define NO_GPU() TORCH_CHECK(false, "oracle/_ref lietorch_backends is a CPU-only build")

torch::Tensor exp_forward_gpu(int, torch::Tensor) { NO_GPU(); return {}; }
std::vector<torch::Tensor> exp_backward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor log_forward_gpu(int, torch::Tensor) { NO_GPU(); return {}; }
std::vector<torch::Tensor> log_backward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor inv_forward_gpu(int, torch::Tensor) { NO_GPU(); return {}; }
std::vector<torch::Tensor> inv_backward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor mul_forward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
std::vector<torch::Tensor> mul_backward_gpu(int, torch::Tensor, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor adj_forward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
std::vector<torch::Tensor> adj_backward_gpu(int, torch::Tensor, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor adjT_forward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
std::vector<torch::Tensor> adjT_backward_gpu(int, torch::Tensor, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor act_forward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
std::vector<torch::Tensor> act_backward_gpu(int, torch::Tensor, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor act4_forward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
std::vector<torch::Tensor> act4_backward_gpu(int, torch::Tensor, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor orthogonal_projector_gpu(int, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor as_matrix_forward_gpu(int, torch::Tensor) { NO_GPU(); return {}; }
torch::Tensor jleft_forward_gpu(int, torch::Tensor, torch::Tensor) { NO_GPU(); return {}; }
