// bwprobe.cu — read-bandwidth probe for the traversal roofline ("L2 bandwidth is measured by the builder with a
// resident-working-set kernel and reported beside" the HBM figure). k_extend reads nodes and triangles with 128-bit loads; on
// C2 the ~60 MB of nodes + leaf triangles stay in the 126 MB L2, so the bandwidth that bounds its bytes there is the L2's, not
// HBM's. vmn_probe_read_bandwidth streams a working set of `bytes` `passes` times with the same access width (LDG.128, .cg =
// cached in L2 only, so L1 hits cannot flatter the number), 8 independent loads in flight per thread, grid = SMs x 8 CTAs of
// 256 threads. A working set well under the L2 size measures L2 read bandwidth (the first pass, which fills it, is run untimed);
// one well above it measures HBM read bandwidth. Diagnostic only: nothing on the render path calls it.
#include "engine.cuh"

namespace vmn {

__device__ __forceinline__ uint4 ld_cg(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

__global__ void __launch_bounds__(256) k_probe_read(const uint4* __restrict__ buf, uint64_t n16, uint32_t passes, uint32_t* __restrict__ sink) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0;
    for (uint32_t p = 0; p < passes; ++p) {
        uint64_t i = tid;
        for (; i + 7 * stride < n16; i += 8 * stride) {
            uint4 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = ld_cg(buf + i + k * stride);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc ^= v[k].x ^ v[k].y ^ v[k].z ^ v[k].w;
        }
        for (; i < n16; i += stride) { const uint4 v = ld_cg(buf + i); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
    }
    if (acc == 0x9E3779B9u) sink[0] = acc; // keeps the loads alive; the buffer is zero-filled, so this never stores
}

} // namespace vmn

extern "C" vmn_status vmn_probe_read_bandwidth(vmn_ctx* c, uint64_t bytes, uint32_t passes, double* gb_per_s) {
    if (!c) return VMN_ERR_INVALID;
    if (!gb_per_s || bytes < (1u << 20) || passes == 0) return vmn::fail(c, VMN_ERR_INVALID, "vmn_probe_read_bandwidth: bytes >= 1 MiB, passes >= 1");
    VMN_CUDA(c, cudaSetDevice(c->device));
    vmn::DevBuf buf, sink;
    vmn_status st = vmn::dev_alloc(c, buf, bytes);
    if (st != VMN_OK) return st;
    st = vmn::dev_alloc(c, sink, 256);
    if (st != VMN_OK) { vmn::dev_free(buf); return st; }
    cudaEvent_t a = nullptr, b = nullptr;
    cudaError_t e = cudaMemsetAsync(buf.p, 0, bytes, c->stream);
    if (e == cudaSuccess) e = cudaEventCreate(&a);
    if (e == cudaSuccess) e = cudaEventCreate(&b);
    float ms = 0.f;
    if (e == cudaSuccess) {
        const uint64_t n16 = bytes / 16;
        const dim3 grid((unsigned)c->sm_count * 8u), block(256);
        vmn::k_probe_read<<<grid, block, 0, c->stream>>>((const uint4*)buf.p, n16, 1u, (uint32_t*)sink.p); // warm: fills L2 / TLBs
        cudaEventRecord(a, c->stream);
        vmn::k_probe_read<<<grid, block, 0, c->stream>>>((const uint4*)buf.p, n16, passes, (uint32_t*)sink.p);
        cudaEventRecord(b, c->stream);
        e = cudaEventSynchronize(b);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, a, b);
        if (e == cudaSuccess) e = cudaGetLastError();
    }
    if (a) cudaEventDestroy(a);
    if (b) cudaEventDestroy(b);
    vmn::dev_free(buf);
    vmn::dev_free(sink);
    if (e != cudaSuccess) return vmn::fail(c, VMN_ERR_CUDA, "vmn_probe_read_bandwidth: %s", cudaGetErrorString(e));
    *gb_per_s = (double)(bytes / 16 * 16) * passes / ((double)ms * 1e6);
    return VMN_OK;
}
