// Symmetric heap on the CUDA virtual-memory API and the NVLS (NVSwitch multicast) data plane (SURVEY §8 N11, N13, N14;
// the reference has no communication code — SURVEY §2.3).
//
// Heap.  Every rank creates its buffers with cuMemCreate (one physical allocation each, exportable as a POSIX file
// descriptor), maps them into its own address space and hands the descriptors to its peers over a unix socket
// (parallel/symm.py); every peer imports and maps them, which gives the same table of W base pointers per buffer that
// the CUDA-IPC heap of comm.cu provides — the peer load/store kernels run on either.  For the buffers that take part
// in collectives a multicast object spanning all W devices is created on top (cuMulticastCreate / AddDevice /
// BindMem) and mapped as one more address range: a store to it lands in all W copies, a `multimem.ld_reduce` from it
// returns the sum over the W copies, both performed by the switch.
//
// Kernels (FTL_COLLECTIVE=nvls; the fixed-order peer-load kernels of comm.cu stay the bitwise-reproducible default):
//   nvls_reduce_scatter_kernel    rank r reads ITS shard of the gradient through the multicast address with
//                                 multimem.ld_reduce (.add, fp32 accumulation in the switch): one 16-byte response
//                                 per vector instead of W-1 peer loads; writes the bf16 sum to its own buffer and the
//                                 sum-of-squares partials for the global norm                          (RS ⊕ norm²)
//   nvls_adamw_allgather_kernel   clip ⊕ AdamW on the owned shard; each updated 16-byte vector is stored ONCE with
//                                 multimem.st and the switch replicates it into all W parameter copies (⊕ AG)
// Per GPU and step this moves 1/W + (W-1)/W of the gradient bytes in the two directions instead of (W-1)/W both ways,
// and the update sends each parameter byte once instead of W-1 times.
//
// Only `multimem.*` instructions and ordinary loads/stores touch multicast or peer memory here.  TMA multicast loads
// here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.
#include <cuda.h>

#include <string.h>
#include <unistd.h>

#include <mutex>

#include "common/util.cuh"

namespace ftl {

// ---- driver entry points (the library does not link libcuda) ---------------------------------------------------------
struct Drv {
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
    CUresult (*MemRelease)(CUmemGenericAllocationHandle);
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
    CUresult (*MemAddressFree)(CUdeviceptr, size_t);
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*MemUnmap)(CUdeviceptr, size_t);
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
    CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
    CUresult (*MemGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
    CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
    CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
    CUresult (*McGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
    CUresult (*DeviceGet)(CUdevice*, int);
    bool ok = false;
};
static Drv g_drv;
static std::once_flag g_drv_once;

template <typename F>
static bool drv_sym(F& fn, const char* name) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
        (void)cudaGetLastError();
        return false;
    }
    fn = reinterpret_cast<F>(p);
    return true;
}

static const Drv& drv() {
    std::call_once(g_drv_once, [] {
        Drv& d = g_drv;
        bool ok = drv_sym(d.MemCreate, "cuMemCreate") && drv_sym(d.MemRelease, "cuMemRelease") &&
                  drv_sym(d.MemAddressReserve, "cuMemAddressReserve") && drv_sym(d.MemAddressFree, "cuMemAddressFree") &&
                  drv_sym(d.MemMap, "cuMemMap") && drv_sym(d.MemUnmap, "cuMemUnmap") && drv_sym(d.MemSetAccess, "cuMemSetAccess") &&
                  drv_sym(d.MemExport, "cuMemExportToShareableHandle") && drv_sym(d.MemImport, "cuMemImportFromShareableHandle") &&
                  drv_sym(d.MemGranularity, "cuMemGetAllocationGranularity") && drv_sym(d.DeviceGetAttribute, "cuDeviceGetAttribute") &&
                  drv_sym(d.DeviceGet, "cuDeviceGet");
        // multicast entry points are optional (older drivers): their absence only disables NVLS
        if (!(drv_sym(d.McCreate, "cuMulticastCreate") && drv_sym(d.McAddDevice, "cuMulticastAddDevice") &&
              drv_sym(d.McBindMem, "cuMulticastBindMem") && drv_sym(d.McGranularity, "cuMulticastGetGranularity")))
            d.McCreate = nullptr;
        d.ok = ok;
    });
    return g_drv;
}

static int cur_device(CUdevice* dev, int* ordinal) {
    int o = 0;
    cudaError_t e = cudaGetDevice(&o);
    if (e != cudaSuccess) return (int)e;
    (void)cudaFree(0);   // make sure the primary context exists
    if (drv().DeviceGet(dev, o) != CUDA_SUCCESS) return (int)cudaErrorInvalidDevice;
    if (ordinal) *ordinal = o;
    return 0;
}

static CUmemAllocationProp local_prop(int ordinal) {
    CUmemAllocationProp p;
    memset(&p, 0, sizeof(p));
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = ordinal;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

// map `handle` (bytes long) somewhere in this process with read/write access for the current device
static int map_rw(CUmemGenericAllocationHandle handle, size_t bytes, size_t align, int ordinal, void** out) {
    const Drv& d = drv();
    CUdeviceptr va = 0;
    CUresult r = d.MemAddressReserve(&va, bytes, align, 0, 0);
    if (r != CUDA_SUCCESS) return 1000 + (int)r;
    r = d.MemMap(va, bytes, 0, handle, 0);
    if (r != CUDA_SUCCESS) { d.MemAddressFree(va, bytes); return 1000 + (int)r; }
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = ordinal;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = d.MemSetAccess(va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) { d.MemUnmap(va, bytes); d.MemAddressFree(va, bytes); return 1000 + (int)r; }
    *out = reinterpret_cast<void*>(va);
    return 0;
}

// ---- kernels ----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bf16x8 mm_ld_reduce_add(const __nv_bfloat16* mc) {
    int4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return *reinterpret_cast<bf16x8*>(&r);
}
__device__ __forceinline__ void mm_st(__nv_bfloat16* mc, const bf16x8& v) {
    const int4& r = *reinterpret_cast<const int4*>(&v);
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w) : "memory");
}

__global__ void __launch_bounds__(512)
nvls_reduce_scatter_kernel(const __nv_bfloat16* __restrict__ mc_g, __nv_bfloat16* __restrict__ g_local, int64_t lo_vec,
                           int64_t nvec, float* __restrict__ partial, const uint32_t* __restrict__ status) {
    __shared__ float red[33];
    float acc = 0.f;
    if (*status == 0) {
        const __nv_bfloat16* src = mc_g + lo_vec * 8;
        __nv_bfloat16* dst = g_local + lo_vec * 8;
        // four independent switch reductions in flight per thread (the round trip through the switch is long)
        const int64_t stride = (int64_t)gridDim.x * blockDim.x;
        for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i0 < nvec; i0 += 4 * stride) {
            bf16x8 s[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (i0 + u * stride < nvec) s[u] = mm_ld_reduce_add(src + (i0 + u * stride) * 8);   // fp32 sum over W copies
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (i0 + u * stride < nvec) {
                    st8(dst + (i0 + u * stride) * 8, s[u]);
                    float f[8];
                    unpack8(s[u], f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc += f[j] * f[j];
                }
            }
        }
    }
    const float total = block_sum(acc, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = total;
}

__global__ void __launch_bounds__(512)
nvls_adamw_allgather_kernel(__nv_bfloat16* __restrict__ mc_p, const __nv_bfloat16* __restrict__ p_local,
                            const __nv_bfloat16* __restrict__ g, __nv_bfloat16* __restrict__ m, __nv_bfloat16* __restrict__ v,
                            int64_t lo_vec, int64_t nvec, const float* __restrict__ hyper, const float* __restrict__ stats,
                            const uint32_t* __restrict__ status, unsigned long long* __restrict__ commits) {
    if (*status != 0 || stats[2] == 0.f) return;   // aborted step or non-finite norm: leave every copy untouched
    if (commits && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(commits, 1ull);
    const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
    const float bc1 = hyper[5], bc2_sqrt = hyper[6], coef = stats[1];
    const float step_size = lr / bc1, decay = 1.f - lr * wd;
    const float vcut = adam_vcut(eps, bc2_sqrt);
    mc_p += lo_vec * 8; p_local += lo_vec * 8; g += lo_vec * 8; m += lo_vec * 8; v += lo_vec * 8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        float pf[8], gf[8], mf[8], vf[8];
        unpack8(ld8_stream(p_local + i * 8), pf);
        unpack8(ld8_stream(g + i * 8), gf);
        unpack8(ld8_stream(m + i * 8), mf);
        unpack8(ld8_stream(v + i * 8), vf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gr = gf[j] * coef;
            mf[j] = mf[j] + (1.f - b1) * (gr - mf[j]);
            vf[j] = b2 * vf[j] + (1.f - b2) * gr * gr;
            const float denom = adam_denom(vf[j], bc2_sqrt, eps, vcut);
            pf[j] = pf[j] * decay - step_size * adam_quot(mf[j], denom);
        }
        st8_stream(m + i * 8, pack8(mf));
        st8_stream(v + i * 8, pack8(vf));
        mm_st(mc_p + i * 8, pack8(pf));            // one store; the switch writes all W parameter copies (own included)
    }
}

}  // namespace ftl

using namespace ftl;

// ---- heap -----------------------------------------------------------------------------------------------------------
// *vmm = 1 when the virtual-memory API with file-descriptor export is usable on the current device,
// *multicast = 1 when the device also supports multicast objects (NVLS).
FTL_API int ftl_vmm_supported(int* vmm, int* multicast) {
    *vmm = 0; *multicast = 0;
    if (!drv().ok) return 0;
    CUdevice dev; int e = cur_device(&dev, nullptr);
    if (e) return e;
    int a = 0;
    if (drv().DeviceGetAttribute(&a, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS) *vmm = a;
    a = 0;
    if (drv().McCreate && drv().DeviceGetAttribute(&a, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS) *multicast = a;
    return 0;
}

// bytes rounded up so that the size suits both a local allocation and a multicast object over `world` devices
FTL_API int64_t ftl_vmm_round(int64_t bytes, int world) {
    if (!drv().ok) return -1;
    int ord = 0; CUdevice dev;
    if (cur_device(&dev, &ord)) return -1;
    CUmemAllocationProp p = local_prop(ord);
    size_t g = 0;
    if (drv().MemGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || g == 0) return -1;
    if (world > 1 && drv().McCreate) {
        CUmulticastObjectProp mp;
        memset(&mp, 0, sizeof(mp));
        mp.numDevices = (unsigned)world; mp.size = (size_t)bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mg = 0;
        if (drv().McGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
    }
    return (int64_t)(((size_t)bytes + g - 1) / g * g);
}

// Allocate `bytes` (already rounded) on the current device, map it, zero it, export a file descriptor for the peers.
FTL_API int ftl_vmm_alloc(int64_t bytes, void** ptr, uint64_t* handle, int* fd) {   // allocates (set-up time only)
    if (!drv().ok) return (int)cudaErrorNotSupported;
    int ord = 0; CUdevice dev; int e = cur_device(&dev, &ord);
    if (e) return e;
    CUmemAllocationProp p = local_prop(ord);
    CUmemGenericAllocationHandle h;
    CUresult r = drv().MemCreate(&h, (size_t)bytes, &p, 0);
    if (r != CUDA_SUCCESS) return 1000 + (int)r;
    if ((e = map_rw(h, (size_t)bytes, 0, ord, ptr))) { drv().MemRelease(h); return e; }
    cudaError_t ce = cudaMemset(*ptr, 0, (size_t)bytes);
    if (ce != cudaSuccess) return (int)ce;
    int out = -1;
    r = drv().MemExport(&out, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) return 1000 + (int)r;
    *handle = (uint64_t)h; *fd = out;
    return 0;
}

// Import a peer's allocation from its file descriptor (received over a unix socket) and map it for the current device.
FTL_API int ftl_vmm_import(int fd, int64_t bytes, void** ptr, uint64_t* handle) {
    if (!drv().ok) return (int)cudaErrorNotSupported;
    int ord = 0; CUdevice dev; int e = cur_device(&dev, &ord);
    if (e) return e;
    CUmemGenericAllocationHandle h;
    CUresult r = drv().MemImport(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) return 1000 + (int)r;
    if ((e = map_rw(h, (size_t)bytes, 0, ord, ptr))) { drv().MemRelease(h); return e; }
    *handle = (uint64_t)h;
    return 0;
}

FTL_API int ftl_vmm_unmap(void* ptr, int64_t bytes, uint64_t handle) {
    if (!drv().ok) return (int)cudaErrorNotSupported;
    CUresult r = CUDA_SUCCESS;
    if (ptr) {
        r = drv().MemUnmap((CUdeviceptr)ptr, (size_t)bytes);
        if (r == CUDA_SUCCESS) r = drv().MemAddressFree((CUdeviceptr)ptr, (size_t)bytes);
    }
    if (handle) drv().MemRelease((CUmemGenericAllocationHandle)handle);
    return r == CUDA_SUCCESS ? 0 : 1000 + (int)r;
}

FTL_API int ftl_fd_close(int fd) { return close(fd); }

// ---- multicast objects -------------------------------------------------------------------------------------------------
FTL_API int ftl_mc_create(int world, int64_t bytes, uint64_t* mc, int* fd) {   // one rank creates, everyone imports
    if (!drv().ok || !drv().McCreate) return (int)cudaErrorNotSupported;
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)world; mp.size = (size_t)bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle h;
    CUresult r = drv().McCreate(&h, &mp);
    if (r != CUDA_SUCCESS) return 1000 + (int)r;
    int out = -1;
    r = drv().MemExport(&out, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) return 1000 + (int)r;
    *mc = (uint64_t)h; *fd = out;
    return 0;
}

FTL_API int ftl_mc_import(int fd, uint64_t* mc) {
    if (!drv().ok || !drv().McCreate) return (int)cudaErrorNotSupported;
    CUmemGenericAllocationHandle h;
    CUresult r = drv().MemImport(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) return 1000 + (int)r;
    *mc = (uint64_t)h;
    return 0;
}

FTL_API int ftl_mc_add_device(uint64_t mc) {   // every rank, before anyone binds
    CUdevice dev; int e = cur_device(&dev, nullptr);
    if (e) return e;
    CUresult r = drv().McAddDevice((CUmemGenericAllocationHandle)mc, dev);
    return r == CUDA_SUCCESS ? 0 : 1000 + (int)r;
}

// bind this rank's physical allocation `mem` as its replica, then map the multicast range for the current device
FTL_API int ftl_mc_bind_map(uint64_t mc, uint64_t mem, int64_t bytes, void** ptr) {
    int ord = 0; CUdevice dev; int e = cur_device(&dev, &ord);
    if (e) return e;
    CUresult r = drv().McBindMem((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem, 0, (size_t)bytes, 0);
    if (r != CUDA_SUCCESS) return 1000 + (int)r;
    return map_rw((CUmemGenericAllocationHandle)mc, (size_t)bytes, 0, ord, ptr);
}

// ---- kernels -----------------------------------------------------------------------------------------------------------
static inline int nvls_grid(int64_t nvec, int max_ctas) {
    int64_t blocks = (nvec + 511) / 512;
    int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)kNumSMs * 4;
    return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

// Reduce shard [lo, lo+n) of the gradient over all ranks through the multicast mapping mc_g; the bf16 sum goes to
// g_local[lo, lo+n) (this rank's own buffer), partial[] (ftl_dp_partials(max_ctas) floats) receives sum-of-squares pieces.
FTL_API int ftl_nvls_reduce_scatter(const void* mc_g, void* g_local, int64_t lo, int64_t n, float* partial,
                                    const void* status, int max_ctas, cudaStream_t stream) {
    if (lo % 8 || n % 8 || n <= 0) return (int)cudaErrorInvalidValue;
    const int grid = max_ctas > 0 ? max_ctas : kNumSMs * 4;   // every partial slot is written (same as ftl_dp_reduce_scatter)
    nvls_reduce_scatter_kernel<<<grid, 512, 0, stream>>>((const __nv_bfloat16*)mc_g, (__nv_bfloat16*)g_local, lo / 8, n / 8,
                                                         partial, (const uint32_t*)status);
    FTL_LAUNCH_CHECK();
    return 0;
}

// clip ⊕ AdamW on [lo, lo+n) ⊕ all-gather: the new parameters are stored through the multicast mapping mc_p.
FTL_API int ftl_nvls_adamw_allgather(void* mc_p, const void* p_local, const void* g, void* m, void* v, int64_t lo, int64_t n,
                                     const float* hyper, const float* stats, const void* status, void* commits, int max_ctas,
                                     cudaStream_t stream) {
    if (lo % 8 || n % 8 || n <= 0) return (int)cudaErrorInvalidValue;
    nvls_adamw_allgather_kernel<<<nvls_grid(n / 8, max_ctas), 512, 0, stream>>>(
        (__nv_bfloat16*)mc_p, (const __nv_bfloat16*)p_local, (const __nv_bfloat16*)g, (__nv_bfloat16*)m, (__nv_bfloat16*)v,
        lo / 8, n / 8, hyper, stats, (const uint32_t*)status, (unsigned long long*)commits);
    FTL_LAUNCH_CHECK();
    return 0;
}
