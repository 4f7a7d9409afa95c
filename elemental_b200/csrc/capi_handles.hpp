// Handle structs behind the opaque C-ABI types (include/elb200.h), shared by capi_engine.cpp and comm.cpp.
#pragma once
#include <memory>

#include "engine.hpp"

struct elb200_grid_s {
    std::shared_ptr<elb::Grid> g;
};
struct elb200_dm_s {
    int dtype;
    // keeps the Grid (its streams) alive until the matrix is gone, whatever order the caller
    // destroys handles in; declared before f/d so that it is released after them
    std::shared_ptr<elb::Grid> keep;
    std::unique_ptr<elb::DistMat<float>> f;
    std::unique_ptr<elb::DistMat<double>> d;
    std::unique_ptr<elb::DistMat<elb::Half>> h;
    elb::Grid* grid() const { return dtype == ELB200_F32 ? f->grid : dtype == ELB200_F64 ? d->grid : h->grid; }
};
