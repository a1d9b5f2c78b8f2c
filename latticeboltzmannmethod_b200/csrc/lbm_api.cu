// lbm_api.cu -- the extern "C" boundary of liblbm_b200.so (include/lbm_b200.h).
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

#include "lbm_internal.h"

namespace lbm {
lbm_status nccl_unique_id(void *out, std::string *why);
}

namespace {
thread_local std::string g_create_error;

bool have_ctx(lbm_ctx *ctx) { return ctx && ctx->model; }

// Every entry point binds the context's device and drops a stale *non-sticky* error another call left behind
// in this thread (ours are read right after each launch; a faulted kernel's error is sticky and survives this),
// so that the next launch check reports its own launch and nothing else.
lbm_status bind_device(lbm_ctx *ctx)
{
    LBM_TRY(ctx->rt.cuda(cudaSetDevice(ctx->rt.device), "cudaSetDevice"));
    ctx->bound = true;
    (void)cudaGetLastError();
    return LBM_OK;
}
}  // namespace

extern "C" {

lbm_status lbm_config_default(lbm_config *c, int model)
{
    if (!c) return LBM_EINVAL;
    std::memset(c, 0, sizeof *c);
    c->model = model;
    c->nz = 1;
    c->slab_count = 1;
    c->tau_phi = 1.0;
    switch (model) {
    case LBM_SHANCHEN_D2Q9:  // shanchen.cpp:11-27
        c->nx = 128; c->ny = 128; c->tau = 1.0; c->g = -5.0;
        break;
    case LBM_BINARY_D2Q9:  // binary.cpp:12-41
        c->nx = 128; c->ny = 128; c->aconst = 0.04; c->kconst = 0.04; c->gammaconst = 1.0; c->wall_gradient = -0.15;
        c->tau_gas = 1.0; c->tau_liq = 1.0; c->tau_phi = 1.0;
        break;
    case LBM_MRT_GS_D2Q9:  // mrt.cpp:8-11,98 with the arguments `10 10`
        c->nx = 40; c->ny = 3; c->omega = 1.0; c->omega_bulk = 1.0; c->force_ampl = 0.002;
        break;
    case LBM_MRT_HERMITE_D2Q9:  // another_mrt.cpp:7-10,52-53
        c->nx = 40; c->ny = 3; c->omega = 0.9; c->omega_bulk = 0.1; c->force_ampl = 0.0;
        break;
    case LBM_TRT_DIFFUSION_D2Q9:  // cylinder.cpp:226-227,33,38,43
        c->nx = 129; c->ny = 129; c->omega = 1.0 / (0.5 + 0.5 / 32.0); c->ce = 1.0 / 3.0; c->conc_wall = 1.0;
        break;
    case LBM_BINARY_SHAPE_D2Q9:  // binary_shape.cpp:17-18,31,37-48
        c->nx = 301; c->ny = 301; c->aconst = 0.04; c->kconst = 0.04; c->gammaconst = 1.0; c->wall_gradient = -0.05;
        c->tau_gas = 1.0; c->tau_liq = 1.0; c->tau_phi = 1.0;
        break;
    case LBM_BINARY_D3Q19:  // QuarterChannel/main.cpp:18-20,48-62
        c->nx = 27; c->ny = 27; c->nz = 52; c->aconst = 0.04; c->kconst = 0.04; c->gammaconst = 1.0;
        c->wall_gradient = 0.0; c->rho_wall = 0.5; c->tau_phi = 2.0; c->tau_liq = 2.5; c->tau_gas = 0.7;
        c->force[0] = 0.0; c->force[1] = 0.0; c->force[2] = 0.00003;
        break;
    default:
        return LBM_EINVAL;
    }
    return LBM_OK;
}

lbm_status lbm_create(const lbm_config *cfg, lbm_ctx **out)
{
    if (!cfg || !out) {
        g_create_error = "lbm_create: null argument";
        return LBM_EINVAL;
    }
    *out = nullptr;
    lbm_ctx *ctx = new (std::nothrow) lbm_ctx;
    if (!ctx) {
        g_create_error = "out of host memory";
        return LBM_ENOMEM;
    }
    ctx->cfg = *cfg;
    ctx->rt.device = cfg->device;
    const char *guard = std::getenv("LBM_GUARD");
    ctx->rt.guard = guard && guard[0] == '1';
    lbm_status st = LBM_OK;
    int ndev = 0;
    st = ctx->rt.cuda(cudaGetDeviceCount(&ndev), "cudaGetDeviceCount");
    if (st == LBM_OK && (cfg->device < 0 || cfg->device >= ndev)) st = ctx->rt.fail(LBM_EINVAL, "no such CUDA device");
    if (st == LBM_OK) st = bind_device(ctx);
    if (st == LBM_OK) {
        st = ctx->rt.cuda(cudaStreamCreateWithFlags(&ctx->rt.stream, cudaStreamNonBlocking), "cudaStreamCreate");
        ctx->rt.own_stream = st == LBM_OK;
    }
    if (st == LBM_OK) {
        switch (cfg->model) {
#ifndef LBM_POP_FLOAT
        case LBM_SHANCHEN_D2Q9: ctx->model = lbm::make_d2q9_shanchen(ctx->rt, ctx->cfg, &st); break;
        case LBM_BINARY_D2Q9: ctx->model = lbm::make_d2q9_binary(ctx->rt, ctx->cfg, &st); break;
        case LBM_MRT_GS_D2Q9:
        case LBM_MRT_HERMITE_D2Q9: ctx->model = lbm::make_d2q9_mrt(ctx->rt, ctx->cfg, &st); break;
        case LBM_TRT_DIFFUSION_D2Q9: ctx->model = lbm::make_d2q9_trt(ctx->rt, ctx->cfg, &st); break;
        case LBM_BINARY_SHAPE_D2Q9: ctx->model = lbm::make_d2q9_binary_shape(ctx->rt, ctx->cfg, &st); break;
#else
        case LBM_SHANCHEN_D2Q9: case LBM_BINARY_D2Q9: case LBM_MRT_GS_D2Q9: case LBM_MRT_HERMITE_D2Q9:
        case LBM_TRT_DIFFUSION_D2Q9: case LBM_BINARY_SHAPE_D2Q9:
            st = ctx->rt.fail(LBM_EINVAL, "liblbm_b200_f32.so holds the D3Q19 model only; the D2Q9 models are in liblbm_b200.so");
            break;
#endif
        case LBM_BINARY_D3Q19: ctx->model = lbm::make_d3q19_binary(ctx->rt, ctx->cfg, &st); break;
        default: st = ctx->rt.fail(LBM_EINVAL, "unknown model");
        }
    }
    if (st != LBM_OK) {
        g_create_error = ctx->rt.err;
        lbm_destroy(ctx);
        return st;
    }
    *out = ctx;
    return LBM_OK;
}

void lbm_destroy(lbm_ctx *ctx)
{
    if (!ctx) return;
    if (ctx->bound) cudaSetDevice(ctx->rt.device);     // a context whose device never existed owns nothing on it
    if (ctx->rt.stream) cudaStreamSynchronize(ctx->rt.stream);
    delete ctx->model;
    ctx->rt.profile_release();
    if (ctx->rt.own_stream && ctx->rt.stream) cudaStreamDestroy(ctx->rt.stream);
    delete ctx;
}

const char *lbm_last_error(const lbm_ctx *ctx) { return ctx ? ctx->rt.err.c_str() : g_create_error.c_str(); }

lbm_status lbm_slab_extent(const lbm_ctx *ctx, int *z_begin, int *nz_local)
{
    if (!ctx || !ctx->model || !z_begin || !nz_local) return LBM_EINVAL;
    return ctx->model->slab_extent(z_begin, nz_local);
}

size_t lbm_local_nodes(const lbm_ctx *ctx) { return ctx && ctx->model ? ctx->model->local_nodes() : 0; }

#define LBM_ENTER(ctx)                                \
    if (!have_ctx(ctx)) return LBM_EINVAL;            \
    LBM_TRY(bind_device(ctx))

lbm_status lbm_set_field(lbm_ctx *ctx, int field, const double *host, size_t n)
{
    LBM_ENTER(ctx);
    if (!host || n != ctx->model->local_nodes()) return ctx->rt.fail(LBM_EINVAL, "lbm_set_field: n must equal lbm_local_nodes()");
    return ctx->model->set_field(field, host);
}

lbm_status lbm_set_geometry(lbm_ctx *ctx, const int *mask, size_t n)
{
    LBM_ENTER(ctx);
    if (ctx->cfg.model != LBM_TRT_DIFFUSION_D2Q9 && ctx->cfg.model != LBM_BINARY_SHAPE_D2Q9) return ctx->rt.fail(LBM_EINVAL, "lbm_set_geometry: this model has no geometry mask");
    if (!mask || n != ctx->model->local_nodes()) return ctx->rt.fail(LBM_EINVAL, "lbm_set_geometry: n must equal lbm_local_nodes()");
    return ctx->model->set_geometry(mask);
}

lbm_status lbm_get_field(lbm_ctx *ctx, int field, double *host, size_t n)
{
    LBM_ENTER(ctx);
    if (!host || n != ctx->model->local_nodes()) return ctx->rt.fail(LBM_EINVAL, "lbm_get_field: n must equal lbm_local_nodes()");
    return ctx->model->get_field(field, host);
}

lbm_status lbm_get_field_async(lbm_ctx *ctx, int field, double *host, size_t n)
{
    LBM_ENTER(ctx);
    if (!host || n != ctx->model->local_nodes()) return ctx->rt.fail(LBM_EINVAL, "lbm_get_field_async: n must equal lbm_local_nodes()");
    return ctx->model->get_field_async(field, host);
}

lbm_status lbm_init(lbm_ctx *ctx)
{
    LBM_ENTER(ctx);
    LBM_TRY(ctx->model->init());
    ctx->initialised = true;
    return LBM_OK;
}

lbm_status lbm_set_populations(lbm_ctx *ctx, int set, const double *host, size_t n)
{
    LBM_ENTER(ctx);
    if (!host || set < 0 || set >= ctx->model->nsets() || n != ctx->model->local_nodes() * ctx->model->q())
        return ctx->rt.fail(LBM_EINVAL, "lbm_set_populations: bad set or n != nodes*Q");
    LBM_TRY(ctx->model->set_populations(set, host));
    ctx->initialised = true;
    return LBM_OK;
}

lbm_status lbm_get_populations(lbm_ctx *ctx, int set, double *host, size_t n)
{
    LBM_ENTER(ctx);
    if (!host || set < 0 || set >= ctx->model->nsets() || n != ctx->model->local_nodes() * ctx->model->q())
        return ctx->rt.fail(LBM_EINVAL, "lbm_get_populations: bad set or n != nodes*Q");
    return ctx->model->get_populations(set, host);
}

lbm_status lbm_step(lbm_ctx *ctx, int nsteps)
{
    LBM_ENTER(ctx);
    if (!ctx->initialised) return ctx->rt.fail(LBM_ESTATE, "lbm_step before lbm_init / lbm_set_populations");
    if (nsteps < 0) return ctx->rt.fail(LBM_EINVAL, "nsteps < 0");
    return ctx->model->step(nsteps);
}

lbm_status lbm_sync(lbm_ctx *ctx)
{
    LBM_ENTER(ctx);
    return ctx->rt.cuda(cudaStreamSynchronize(ctx->rt.stream), "cudaStreamSynchronize");
}

lbm_status lbm_diagnostics(lbm_ctx *ctx, double *mass, double momentum[3], double *phase_mass, int *has_nan)
{
    LBM_ENTER(ctx);
    return ctx->model->diagnostics(mass, momentum, phase_mass, has_nan);
}

void *lbm_get_stream(lbm_ctx *ctx) { return ctx ? (void *)ctx->rt.stream : nullptr; }

lbm_status lbm_set_stream(lbm_ctx *ctx, void *stream)
{
    LBM_ENTER(ctx);
    LBM_CUDA(ctx->rt, cudaStreamSynchronize(ctx->rt.stream));
    if (ctx->rt.own_stream) cudaStreamDestroy(ctx->rt.stream);
    ctx->rt.stream = (cudaStream_t)stream;
    ctx->rt.own_stream = false;
    return LBM_OK;
}

unsigned long long lbm_launch_count(const lbm_ctx *ctx) { return ctx ? ctx->rt.launches : 0; }
unsigned long long lbm_device_bytes(const lbm_ctx *ctx)
{
    unsigned long long n = 0;
    if (ctx)
        for (const auto &kv : ctx->rt.allocs) n += kv.second.bytes;
    return n;
}
int lbm_population_bytes(const lbm_ctx *ctx) { return ctx && ctx->model ? ctx->model->pop_bytes() : 0; }
unsigned long long lbm_guard_check(lbm_ctx *ctx)
{
    if (!have_ctx(ctx) || bind_device(ctx) != LBM_OK) return 0;
    return ctx->rt.guard_check();
}
unsigned long long lbm_phase_fallbacks(lbm_ctx *ctx)
{
    if (!have_ctx(ctx) || bind_device(ctx) != LBM_OK) return 0;
    return ctx->model->fallbacks();
}

lbm_status lbm_profile(lbm_ctx *ctx, int enable)
{
    LBM_ENTER(ctx);
    ctx->rt.profile = enable != 0;
    if (!enable) ctx->rt.profile_release();   // also drops a record nobody asked the report of
    return LBM_OK;
}

lbm_status lbm_profile_report(lbm_ctx *ctx, char *buf, size_t buf_bytes)
{
    LBM_ENTER(ctx);
    if (!buf || buf_bytes < 2) return LBM_EINVAL;
    std::string js = "{";
    bool first = true;
    for (auto &kv : ctx->rt.profile_collect()) {
        char line[256];
        snprintf(line, sizeof line, "%s\"%s\": {\"launches\": %d, \"ms\": %.6f}", first ? "" : ", ", kv.first.c_str(),
                 kv.second.first, kv.second.second);
        js += line;
        first = false;
    }
    js += "}";
    if (js.size() + 1 > buf_bytes) return ctx->rt.fail(LBM_EINVAL, "lbm_profile_report: buffer too small");
    std::memcpy(buf, js.c_str(), js.size() + 1);
    return LBM_OK;
}

void *lbm_host_alloc(size_t bytes)
{
    void *p = nullptr;
    return cudaMallocHost(&p, bytes) == cudaSuccess ? p : nullptr;
}
void lbm_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

lbm_status lbm_comm_unique_id(void *id_out)
{
    if (!id_out) return LBM_EINVAL;
    return lbm::nccl_unique_id(id_out, &g_create_error);
}

lbm_status lbm_comm_attach(lbm_ctx *ctx, const void *id)
{
    LBM_ENTER(ctx);
    if (!id) return ctx->rt.fail(LBM_EINVAL, "null communicator id");
    return ctx->model->comm_attach(id);
}

lbm_status lbm_slab_halo(lbm_ctx *ctx, int halo_id, void **device_ptr, size_t *count)
{
    LBM_ENTER(ctx);
    if (!device_ptr || !count) return LBM_EINVAL;
    return ctx->model->slab_halo(halo_id, device_ptr, count);
}
lbm_status lbm_slab_init_fields(lbm_ctx *ctx) { LBM_ENTER(ctx); return ctx->model->slab_init_fields(); }
lbm_status lbm_slab_init_pops(lbm_ctx *ctx) { LBM_ENTER(ctx); return ctx->model->slab_init_pops(); }
lbm_status lbm_slab_after_init(lbm_ctx *ctx)
{
    LBM_ENTER(ctx);
    LBM_TRY(ctx->model->slab_after_init());
    ctx->initialised = true;
    return LBM_OK;
}
lbm_status lbm_slab_collide(lbm_ctx *ctx, int write_fields) { LBM_ENTER(ctx); return ctx->model->slab_collide(write_fields); }
lbm_status lbm_slab_finish(lbm_ctx *ctx) { LBM_ENTER(ctx); return ctx->model->slab_finish(); }

}  // extern "C"
