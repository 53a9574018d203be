// Scratch: does programmatic dependent launch survive stream capture, and what does it save per node?
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); return 1; } } while (0)
__global__ void __launch_bounds__(256) k(int* p, int n, int pdl) {
    if (pdl) {
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;");
    }
    if (p[0] == 12345) return;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i + 1] += 1;
}
template <typename... A>
cudaError_t launch(void (*f)(A...), dim3 g, dim3 b, cudaStream_t s, bool pdl, A... a) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = g, cfg.blockDim = b, cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at, cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, f, a...);
}
int main() {
    int* d; const int n = 1 << 20;
    CK(cudaMalloc(&d, (n + 1) * 4)); CK(cudaMemset(d, 0, (n + 1) * 4));
    cudaStream_t s, aux; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&aux, cudaStreamNonBlocking));
    cudaEvent_t e0, e1, f0, f1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventCreateWithFlags(&f0, cudaEventDisableTiming); cudaEventCreateWithFlags(&f1, cudaEventDisableTiming);
    for (int work = 0; work < 2; work++)
    for (int pdl = 0; pdl < 2; pdl++) {
        for (int fork = 0; fork < 2; fork++) {
            cudaGraph_t g; cudaGraphExec_t ge;
            CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
            for (int i = 0; i < 24; i++) {
                if (fork && i == 10) {
                    CK(cudaEventRecord(f0, s)); CK(cudaStreamWaitEvent(aux, f0, 0));
                    CK(launch(k, dim3(64), dim3(256), aux, pdl, d, 0, pdl));
                    CK(launch(k, dim3(64), dim3(256), aux, pdl, d, 0, pdl));
                }
                CK(launch(k, dim3(1184), dim3(256), s, pdl, d, work ? n : 0, pdl));
                if (fork && i == 12) { CK(cudaEventRecord(f1, aux)); CK(cudaStreamWaitEvent(s, f1, 0)); }
            }
            CK(cudaStreamEndCapture(s, &g));
            CK(cudaGraphInstantiate(&ge, g, 0));
            for (int w = 0; w < 20; w++) CK(cudaGraphLaunch(ge, s));
            CK(cudaEventRecord(e0, s));
            for (int w = 0; w < 200; w++) CK(cudaGraphLaunch(ge, s));
            CK(cudaEventRecord(e1, s)); CK(cudaStreamSynchronize(s));
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            printf("work=%d pdl=%d fork=%d: %.2f us per graph of 24(+2) kernels, %.2f us per node\n", work, pdl, fork, ms * 1e3 / 200, ms * 1e3 / 200 / 24);
            cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
        }
    }
    int h[4]; CK(cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost)); printf("check %d %d\n", h[1], h[2]);
    return 0;
}
