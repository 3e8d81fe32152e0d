// Feasibility probe: a CUDA-graph WHILE node whose condition is set from device code,
// plus timing of body iterations (what the fold loop would become).
#include <cuda_runtime.h>
#include <stdio.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
__global__ void body(int *ctr, int *work, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) work[i] += 1;
}
__global__ void advance(int *ctr, cudaGraphConditionalHandle h, int limit) {
  int v = ++ctr[0];
  cudaGraphSetConditional(h, v < limit ? 1u : 0u);
}
int main() {
  int *ctr, *work; const int n = 1 << 20;
  CK(cudaMalloc(&ctr, 64)); CK(cudaMalloc(&work, n * 4));
  CK(cudaMemset(ctr, 0, 64)); CK(cudaMemset(work, 0, n * 4));
  cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  cudaGraph_t g; CK(cudaGraphCreate(&g, 0));
  cudaGraphConditionalHandle h;
  CK(cudaGraphConditionalHandleCreate(&h, g, 1, cudaGraphCondAssignDefault));
  cudaGraphNodeParams cp = {}; cp.type = cudaGraphNodeTypeConditional;
  cp.conditional.handle = h; cp.conditional.type = cudaGraphCondTypeWhile; cp.conditional.size = 1;
  cudaGraphNode_t cn; CK(cudaGraphAddNode(&cn, g, nullptr, 0, &cp));
  cudaGraph_t bg = cp.conditional.phGraph_out[0];
  // body: capture a stream into the body graph
  cudaStream_t cs; CK(cudaStreamCreate(&cs));
  CK(cudaStreamBeginCaptureToGraph(cs, bg, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
  for (int k = 0; k < 5; k++) body<<<n / 256, 256, 0, cs>>>(ctr, work, n);
  advance<<<1, 1, 0, cs>>>(ctr, h, 13);
  CK(cudaStreamEndCapture(cs, nullptr));
  cudaGraphExec_t ge; CK(cudaGraphInstantiate(&ge, g, 0));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 3; rep++) {
    CK(cudaMemsetAsync(ctr, 0, 64, s));
    CK(cudaEventRecord(e0, s));
    CK(cudaGraphLaunch(ge, s));
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int hc, hw; cudaMemcpy(&hc, ctr, 4, cudaMemcpyDeviceToHost); cudaMemcpy(&hw, work, 4, cudaMemcpyDeviceToHost);
    printf("rep %d: iterations %d work[0] %d  %.3f ms (%.1f us / iteration of 6 kernels)\n", rep, hc, hw, ms, ms * 1000 / hc);
  }
  // the same as stream launches with a host sync per iteration
  int *hp; CK(cudaHostAlloc(&hp, 64, 0));
  for (int rep = 0; rep < 3; rep++) {
    CK(cudaMemsetAsync(ctr, 0, 64, s));
    CK(cudaEventRecord(e0, s));
    for (int it = 0; it < 13; it++) {
      for (int k = 0; k < 5; k++) body<<<n / 256, 256, 0, s>>>(ctr, work, n);
      CK(cudaMemcpyAsync(hp, ctr, 4, cudaMemcpyDeviceToHost, s));
      CK(cudaStreamSynchronize(s));
    }
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("stream+sync rep %d: %.3f ms (%.1f us / iteration)\n", rep, ms, ms * 1000 / 13);
  }
  printf("CONDGRAPH OK\n");
  return 0;
}
