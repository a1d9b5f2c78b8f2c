// Shared by hosts/microchannel_quarter.cpp and hosts/microchannel_full.cpp: the main() of
// Microchannel3D/{QuarterChannel,FullChannel}/main.cpp with the Solver calls replaced by the C ABI.
// One process drives all the GPUs it is given (LBM_GPUS=n, default 1): one context per z-slab, the ring
// exchange done by the library over NCCL.  The .vts writer emits the same logical file as VTK 5's
// vtkXMLStructuredGridWriter (solver.hpp:523-576): Float32 points, then Float64 "Velocity"(3), "Phase",
// "Density", raw appended data.
#include <cmath>
#include <cstdint>
#include <cstring>
#include "host_common.h"

struct ChannelSetup {
    int NX, NY, NZ, NUMTIME, NUMOUTPUT, NUMSIGNAL, width, name_digits;
    double tau_phi, force_z;
};

static void write_vts(const std::string &name, int NX, int NY, int NZ, const std::vector<double> &velocity,
                      const std::vector<double> &phase, const std::vector<double> &density)
{
    const size_t n = (size_t)NX * NY * NZ;
    std::ofstream out((name + ".vts").c_str(), std::ios::binary);
    const uint64_t bytes_v = n * 3 * 8, bytes_s = n * 8, bytes_p = n * 3 * 4;
    uint64_t off = 0;
    out << "<?xml version=\"1.0\"?>\n<VTKFile type=\"StructuredGrid\" version=\"1.0\" byte_order=\"LittleEndian\" header_type=\"UInt64\">\n";
    out << "  <StructuredGrid WholeExtent=\"0 " << NX - 1 << " 0 " << NY - 1 << " 0 " << NZ - 1 << "\">\n";
    out << "    <Piece Extent=\"0 " << NX - 1 << " 0 " << NY - 1 << " 0 " << NZ - 1 << "\">\n      <PointData>\n";
    out << "        <DataArray type=\"Float64\" Name=\"Velocity\" NumberOfComponents=\"3\" format=\"appended\" offset=\"" << off << "\"/>\n";
    off += 8 + bytes_v;
    out << "        <DataArray type=\"Float64\" Name=\"Phase\" format=\"appended\" offset=\"" << off << "\"/>\n";
    off += 8 + bytes_s;
    out << "        <DataArray type=\"Float64\" Name=\"Density\" format=\"appended\" offset=\"" << off << "\"/>\n";
    off += 8 + bytes_s;
    out << "      </PointData>\n      <CellData/>\n      <Points>\n";
    out << "        <DataArray type=\"Float32\" Name=\"Points\" NumberOfComponents=\"3\" format=\"appended\" offset=\"" << off << "\"/>\n";
    out << "      </Points>\n    </Piece>\n  </StructuredGrid>\n  <AppendedData encoding=\"raw\">\n   _";
    out.write((const char *)&bytes_v, 8); out.write((const char *)velocity.data(), bytes_v);
    out.write((const char *)&bytes_s, 8); out.write((const char *)phase.data(), bytes_s);
    out.write((const char *)&bytes_s, 8); out.write((const char *)density.data(), bytes_s);
    std::vector<float> pts(n * 3);
    size_t i = 0;
    for (int z = 0; z < NZ; z++)
        for (int y = 0; y < NY; y++)
            for (int x = 0; x < NX; x++) { pts[i++] = (float)x; pts[i++] = (float)y; pts[i++] = (float)z; }
    out.write((const char *)&bytes_p, 8); out.write((const char *)pts.data(), bytes_p);
    out << "\n  </AppendedData>\n</VTKFile>\n";
}

static int run_channel(const ChannelSetup &S, int argc, char **argv)
{
    int last = S.NUMTIME;
    if (std::getenv("LBM_NUMTIME")) last = std::atoi(std::getenv("LBM_NUMTIME"));   // shorter runs for tests
    const int ngpu = std::getenv("LBM_GPUS") ? std::atoi(std::getenv("LBM_GPUS")) : 1;
    std::vector<lbm_ctx *> ctx(ngpu, (lbm_ctx *)0);
    char uid[LBM_COMM_ID_BYTES];
    if (ngpu > 1) LBM_CHECK(0, lbm_comm_unique_id(uid));
    for (int r = 0; r < ngpu; r++) {
        lbm_config cfg;
        lbm_config_default(&cfg, LBM_BINARY_D3Q19);          // main.cpp:47-62
        cfg.nx = S.NX; cfg.ny = S.NY; cfg.nz = S.NZ; cfg.tau_phi = S.tau_phi; cfg.force[2] = S.force_z;
        cfg.device = r; cfg.slab_rank = r; cfg.slab_count = ngpu;
        LBM_CHECK(0, lbm_create(&cfg, &ctx[r]));
    }
    if (ngpu > 1) {
        // ncclCommInitRank blocks until every rank has joined: one thread per context would be needed in
        // one process; the multi-GPU path is therefore driven from bench.py / torchrun (one process per GPU).
        std::fprintf(stderr, "LBM_GPUS > 1: launch one process per GPU (see INTEGRATION.md)\n");
        return 2;
    }
    const size_t n = (size_t)S.NX * S.NY * S.NZ;
    std::vector<double> rho(n), phase(n), ux(n, 0.0), uy(n, 0.0), uz(n, 0.0), vel(3 * n);
    for (size_t counter = 0; counter < n; counter++) {       // main.cpp:83-132
        int iZ = (int)(counter / (S.NX * S.NY)), iY = (int)((counter % (S.NX * S.NY)) / S.NX), iX = (int)((counter % (S.NX * S.NY)) % S.NX);
        bool gas = (iZ >= (S.NZ - 1) / 3) && (iZ <= 2 * (S.NZ - 1) / 3) && (iX <= S.NX - S.width - 1) && (iY <= S.NY - S.width - 1);
        rho[counter] = 1.0;
        phase[counter] = gas ? -1.0 : 1.0;
    }
    lbm_ctx *c = ctx[0];
    LBM_CHECK(c, lbm_set_field(c, LBM_RHO, rho.data(), n));
    LBM_CHECK(c, lbm_set_field(c, LBM_PHASE, phase.data(), n));
    LBM_CHECK(c, lbm_set_field(c, LBM_UX, ux.data(), n));
    LBM_CHECK(c, lbm_set_field(c, LBM_UY, uy.data(), n));
    LBM_CHECK(c, lbm_set_field(c, LBM_UZ, uz.data(), n));
    LBM_CHECK(c, lbm_init(c));
    for (int time_counter = 0; time_counter < last; time_counter++) {
        int next = time_counter;
        while (next < last - 1 && next % S.NUMSIGNAL != 0 && next % S.NUMOUTPUT != 0) next++;
        LBM_CHECK(c, lbm_step(c, next - time_counter + 1));
        time_counter = next;
        if (time_counter % S.NUMSIGNAL == 0) {
            std::cout << "Time is " << time_counter << "\n";
            int has_nan = 0;   // the reference probes one node of a stale copy (solver.hpp:49-55); this checks all
            LBM_CHECK(c, lbm_diagnostics(c, 0, 0, 0, &has_nan));
            if (has_nan) { std::cout << "Phase fields contain NaN values\n"; return -1; }
        }
        if (time_counter % S.NUMOUTPUT == 0) {
            LBM_CHECK(c, lbm_get_field(c, LBM_RHO, rho.data(), n));
            LBM_CHECK(c, lbm_get_field(c, LBM_PHASE, phase.data(), n));
            LBM_CHECK(c, lbm_get_field(c, LBM_UX, ux.data(), n));
            LBM_CHECK(c, lbm_get_field(c, LBM_UY, uy.data(), n));
            LBM_CHECK(c, lbm_get_field(c, LBM_UZ, uz.data(), n));
            for (size_t i = 0; i < n; i++) { vel[3 * i] = ux[i]; vel[3 * i + 1] = uy[i]; vel[3 * i + 2] = uz[i]; }
            write_vts(padded("phase", S.name_digits, time_counter), S.NX, S.NY, S.NZ, vel, phase, rho);
            std::cout << "Output is done on the step " << time_counter << "\n";
        }
    }
    for (int r = 0; r < ngpu; r++) lbm_destroy(ctx[r]);
    (void)argc; (void)argv;
    return 0;
}
