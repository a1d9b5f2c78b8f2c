// Shared by hosts/microchannel_quarter.cpp and hosts/microchannel_full.cpp: the main() of
// Microchannel3D/{QuarterChannel,FullChannel}/main.cpp (:40-160) against hosts/solver_b200.h, whose Solver has the
// reference Solver<D>'s method names on top of the C ABI.  LBM_NUMTIME shortens the run (tests).  Extensions:
// LBM_RESTART=<name> continues bit-exactly from <name>.pop, LBM_CHECKPOINT=<name> writes one at the end.
#include "solver_b200.h"

struct ChannelSetup {
    int NX, NY, NZ, NUMTIME, NUMOUTPUT, NUMSIGNAL, width, name_digits;
    double tau_phi, force_z;
};

static int run_channel(const ChannelSetup &S, int argc, char **argv)
{
    using std::cout;
    const int NX = S.NX, NY = S.NY, NZ = S.NZ, NUM = NX * NY * NZ, width = S.width;
    int NUMTIME = S.NUMTIME;
    if (std::getenv("LBM_NUMTIME")) NUMTIME = std::atoi(std::getenv("LBM_NUMTIME"));
    const double rhol = 1.0, rhog = 1.0;

    // the parameter list of main.cpp:48-63, same names and values
    ParamsList params;
    const struct { const char *name; double value; } table[] = {
        {"omega", 1.0}, {"aconst", 0.04}, {"kconst", 0.04}, {"gammaconst", 1.0}, {"phase_gradient", 0.0}, {"phase_wall", 0.0},
        {"rho_wall", 0.5}, {"tau_phi", S.tau_phi}, {"tau_liq", 2.5}, {"tau_gas", 0.7}, {"force_x", 0.0}, {"force_y", 0.0},
        {"force_z", S.force_z}, {"NX", (double)NX}, {"NY", (double)NY}, {"NZ", (double)NZ}};
    for (size_t i = 0; i < sizeof table / sizeof table[0]; i++) params.add(table[i].name, table[i].value);

    Solver solver(params);

    double rho_temp, phase_temp, u_temp[3] = {0.0, 0.0, 0.0};
    for (int counter = 0; counter < NUM; counter++) {
        int iZ = counter / (NX * NY);
        int iY = (counter % (NX * NY)) / NX;
        int iX = (counter % (NX * NY)) % NX;
        if ((iZ >= (NZ - 1) / 3) && (iZ <= 2 * (NZ - 1) / 3) && (iX <= NX - width - 1) && (iY <= NY - width - 1)) {
            rho_temp = rhog;
            phase_temp = -1.0;
        } else {
            rho_temp = rhol;
            phase_temp = 1.0;
        }
        solver.putDensity(iX, iY, iZ, iX, iY, iZ, rho_temp);
        solver.putPhase(iX, iY, iZ, iX, iY, iZ, phase_temp);
        solver.putVelocity(iX, iY, iZ, iX, iY, iZ, u_temp);
    }

    int first = 0;
    if (std::getenv("LBM_RESTART"))
        first = solver.loadPopulations(std::getenv("LBM_RESTART")) + 1;
    else
        solver.init();

    for (int time_counter = first; time_counter < NUMTIME; time_counter++) {
        // the steps up to the next one that prints or writes go to the device as one batch
        int next = time_counter;
        while (next < NUMTIME - 1 && next % S.NUMSIGNAL != 0 && next % S.NUMOUTPUT != 0) next++;
        solver.collide_stream(next - time_counter + 1);
        time_counter = next;

        if (time_counter % S.NUMSIGNAL == 0) {
            cout << "Time is " << time_counter << "\n";
            if (solver.checkNAN()) {
                cout << "Phase fields contain NaN values\n";
                return -1;
            }
        }
        if (time_counter % S.NUMOUTPUT == 0) {
            solver.writeWholeDensityPhaseVelocity(padded("phase", S.name_digits, time_counter));
            cout << "Output is done on the step " << time_counter << "\n";
        }
    }
    if (std::getenv("LBM_CHECKPOINT")) solver.savePopulations(std::getenv("LBM_CHECKPOINT"), NUMTIME - 1);
    (void)argc; (void)argv;
    return 0;
}
