// Launcher of the persistent-thread path kernel (persist.cuh); its own translation unit so that the eight
// instantiations compile beside rptx.cu / photon.cu.
#define RPTX_NOINLINE_MATH 1  // one copy of f64 division / square root for the whole-path kernel (common.cuh)
#include "persist.cuh"

#include "host_types.hpp"

namespace rptx_host {

using namespace rptx;

namespace {
typedef void (*PersistKernel)(const SceneDev, const MediumDev, const PersistArgs, unsigned long long*);

PersistKernel pick(bool medium, bool probe, bool small) {
    if (medium) {
        if (probe) return small ? k_path_persist<true, true, true> : k_path_persist<true, true, false>;
        return small ? k_path_persist<true, false, true> : k_path_persist<true, false, false>;
    }
    if (probe) return small ? k_path_persist<false, true, true> : k_path_persist<false, true, false>;
    return small ? k_path_persist<false, false, true> : k_path_persist<false, false, false>;
}
}  // namespace

// resident blocks of the kernel on the whole device: the grid of the persistent launch
int persist_grid(int sm_count, bool medium, bool probe, bool small) {
    int per_sm = 0;
    PersistKernel k = pick(medium, probe, small);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PersistShared));
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kPersistBlock, sizeof(PersistShared)) != cudaSuccess || per_sm < 1) per_sm = 1;
    return sm_count * per_sm;
}

cudaError_t launch_path_persist(bool medium, bool probe, bool small, int blocks, const SceneDev& sc, const MediumDev& md, const PersistArgs& a,
                                unsigned long long* stats, cudaStream_t st) {
    PersistKernel k = pick(medium, probe, small);
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PersistShared));
    if (e != cudaSuccess) return e;
    k<<<blocks, kPersistBlock, sizeof(PersistShared), st>>>(sc, md, a, stats);
    return cudaGetLastError();
}

}  // namespace rptx_host
