// Exercises every public method of hosts/solver_b200.h on a small channel and checks the two restart paths:
//   macroscopic (load_file, solver.hpp:400-520): fields written to a .vts come back identical;
//   exact (savePopulations / loadPopulations): a run continued from the dump equals the uninterrupted run bit for bit.
// Writes its files into the current directory; prints "solver_selftest OK" on success.  Run by tests/test_gpu_hosts.py,
// which also compares the text files with the fields read through the ctypes binding.
#include "solver_b200.h"

static ParamsList params_for(int NX, int NY, int NZ)
{
    ParamsList p;
    p.add("omega", 1.0); p.add("aconst", 0.04); p.add("kconst", 0.04); p.add("gammaconst", 1.0);
    p.add("phase_gradient", 0.02); p.add("phase_wall", 0.0); p.add("rho_wall", 0.5);
    p.add("tau_phi", 2.0); p.add("tau_liq", 2.5); p.add("tau_gas", 0.7);
    p.add("force_x", 0.0); p.add("force_y", 0.0); p.add("force_z", 0.00003);
    p.add("NX", NX); p.add("NY", NY); p.add("NZ", NZ);
    return p;
}

static void bubble(Solver &s, int width)
{
    double u[3] = {0.0, 0.0, 0.0};
    s.putDensity(1.0);
    s.putPhase(1.0);
    s.putVelocity(u);
    s.putPhase(0, 0, (s.NZ - 1) / 3, s.NX - width - 1, s.NY - width - 1, 2 * (s.NZ - 1) / 3, -1.0);
}

static bool same(const std::vector<double> &a, const std::vector<double> &b)
{
    return a.size() == b.size() && std::memcmp(a.data(), b.data(), a.size() * sizeof(double)) == 0;
}

int main()
{
    const int NX = 14, NY = 11, NZ = 16;
    Solver a(params_for(NX, NY, NZ));
    bubble(a, 4);
    a.init();
    a.collide_stream();
    a.collide_stream(6);
    if (a.checkNAN()) { std::printf("NaN\n"); return 1; }
    a.writeWholeDensityPhaseVelocity("all");
    a.writeWholeDensity("density"); a.writeWholePhase("phase"); a.writeWholeVelocity("velocity");
    a.writeTextXZVelocity("vel_xz"); a.writeTextXYVelocity("vel_xy"); a.writeTextYZVelocity("vel_yz");
    a.writeTextWholeVelocity("vel_whole"); a.writeTextWholePhase("phase_whole"); a.writeTextWholeDensity("density_whole");
    a.writeTextXYPhase("phase_xy"); a.writeTextXZPhase("phase_xz"); a.writeTextYZPhase("phase_yz");
    a.writeTextXYDensity("density_xy"); a.writeTextXZDensity("density_xz"); a.writeTextYZDensity("density_yz");
    a.writeVTKWholePhase("phase_legacy");
    a.writeLocalDensity("local_density"); a.writeLocalPhase("local_phase");
    a.writeLocalTextDensity("local_density_text"); a.writeLocalTextPhase("local_phase_text");

    // macroscopic restart: a solver started from the file equals a solver given the same fields node by node
    // (init() itself is not the identity on the fields: solver.hpp:877 recomputes them from the new populations)
    Solver b(params_for(NX, NY, NZ));
    b.load_file("all");
    b.init();
    b.getWholeDensity(); b.getWholePhase(); b.getWholeVelocity();
    Solver c(params_for(NX, NY, NZ));
    for (int z = 0; z < NZ; z++)
        for (int y = 0; y < NY; y++)
            for (int x = 0; x < NX; x++) {
                const size_t i = ((size_t)z * NY + y) * NX + x;
                c.putDensity(x, y, z, x, y, z, a.density[i]);
                c.putPhase(x, y, z, x, y, z, a.phase[i]);
                c.putVelocity(x, y, z, x, y, z, &a.velocity[3 * i]);
            }
    c.init();
    c.getWholeDensity(); c.getWholePhase(); c.getWholeVelocity();
    if (!same(b.density, c.density) || !same(b.phase, c.phase) || !same(b.velocity, c.velocity)) { std::printf("load_file mismatch\n"); return 2; }

    // exact restart
    a.savePopulations("dump", 6);
    a.collide_stream(9);
    a.getWholeDensity(); a.getWholePhase(); a.getWholeVelocity();
    Solver d(params_for(NX, NY, NZ));
    bubble(d, 4);
    const int t = d.loadPopulations("dump");
    d.collide_stream(9);
    d.getWholeDensity(); d.getWholePhase(); d.getWholeVelocity();
    if (t != 6 || !same(a.density, d.density) || !same(a.phase, d.phase) || !same(a.velocity, d.velocity)) { std::printf("exact restart mismatch\n"); return 3; }
    a.writeWholeDensityPhaseVelocity("after_restart");
    std::printf("solver_selftest OK\n");
    return 0;
}
