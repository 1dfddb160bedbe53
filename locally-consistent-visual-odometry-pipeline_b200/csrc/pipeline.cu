// Pipeline-level state and entry points (filled in below).
#include "ctx.cuh"

namespace vo {

struct Pipeline {
    int track_capacity = 0, hyp_capacity = 0;
};

int pipeline_init(Ctx* c, int track_capacity, int hyp_capacity)
{
    c->pipe = new (std::nothrow) Pipeline();
    if (!c->pipe) return set_error(c, VO_ERR_NOMEM, "out of host memory");
    c->pipe->track_capacity = track_capacity;
    c->pipe->hyp_capacity = hyp_capacity;
    return VO_OK;
}

void pipeline_free(Ctx* c)
{
    delete c->pipe;
    c->pipe = nullptr;
}

}  // namespace vo
