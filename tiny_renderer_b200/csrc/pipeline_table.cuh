// pipeline_table.cuh — THE list of shader pipelines (ShaderPipeline::new, /root/reference/src/scene/shader.rs:97-112).
// Adding a pipeline = one header under pipelines/ + one X(...) entry and one #include below.  Everything else — the ids,
// the name lookup, k_vertex_setup<PIPE> / k_raster<PIPE, THREADS> instantiations, the host dispatch table in tinyr.cu —
// is generated from TR_PIPELINES.
#pragma once
#include "tr_prepare.h"

#define TR_PIPELINES(X) X(Default) X(Phong) X(NormalMap) X(Specular) X(Darboux) X(Shadow) X(Occlusion)

namespace tr {
enum Pipeline : int {
#define TR_X(N) kPipe##N,
    TR_PIPELINES(TR_X)
#undef TR_X
    kPipeCount
};
}  // namespace tr

#include "pipelines/default.cuh"
#include "pipelines/phong.cuh"
#include "pipelines/normal_map.cuh"
#include "pipelines/specular.cuh"
#include "pipelines/darboux.cuh"
#include "pipelines/shadow.cuh"
#include "pipelines/occlusion.cuh"
