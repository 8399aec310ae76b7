#include <cuda_runtime.h>
#include <cstdio>
__global__ void body(int* c, cudaGraphConditionalHandle h) { int v = ++*c; cudaGraphSetConditional(h, v < 5 ? 1u : 0u); }
int main() {
  cudaGraph_t g; cudaGraphCreate(&g, 0);
  cudaGraphConditionalHandle h; cudaGraphConditionalHandleCreate(&h, g, 1, cudaGraphCondAssignDefault);
  cudaGraphNodeParams p = {}; p.type = cudaGraphNodeTypeConditional; p.conditional.handle = h; p.conditional.type = cudaGraphCondTypeWhile; p.conditional.size = 1;
  cudaGraphNode_t n; cudaError_t e = cudaGraphAddNode(&n, g, nullptr, 0, &p); printf("%s\n", cudaGetErrorString(e));
  cudaGraph_t b = p.conditional.phGraph_out[0];
  int* c; cudaMalloc(&c, 4); cudaMemset(c, 0, 4);
  cudaKernelNodeParams kp = {}; void* args[2] = {&c, &h}; kp.func = (void*)body; kp.gridDim = dim3(1); kp.blockDim = dim3(1); kp.kernelParams = args;
  cudaGraphNode_t kn; e = cudaGraphAddKernelNode(&kn, b, nullptr, 0, &kp); printf("%s\n", cudaGetErrorString(e));
  cudaGraphExec_t x; e = cudaGraphInstantiate(&x, g, 0); printf("%s\n", cudaGetErrorString(e));
  cudaGraphLaunch(x, 0); cudaDeviceSynchronize(); int hc; cudaMemcpy(&hc, c, 4, cudaMemcpyDeviceToHost); printf("count %d\n", hc);
}
