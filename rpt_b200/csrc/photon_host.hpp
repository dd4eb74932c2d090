// Host-side object behind rptx_photons (PhotonList, src/photon.rs:141-171), shared by photon.cu and comm.cu.
#pragma once
#include "host_types.hpp"

namespace rptx {
struct PhotonRec;  // photon.cuh
}

struct rptx_photons {
    rptx_ctx* ctx = nullptr;
    rptx::PhotonRec* d_rows[2] = {nullptr, nullptr};  // [0] surface, [1] volume
    uint64_t n[2] = {0, 0}, cap[2] = {0, 0};
    uint64_t emitted = 0, rays = 0;
    double emit_ms = 0;
};

namespace rptx_host {
// make list `which` hold exactly n rows (contents of new rows undefined); photon.cu
int photons_resize(rptx_photons* p, int which, uint64_t n);
}
