// trace_bwd.cuh -- adjoint kernel (filled in by the backward milestone)
#pragma once
#include "trace_geom.cuh"
