// Drop-in for Microchannel3D/FullChannel/main.cpp (:19-33: 92x52x1500, 250001 steps, output every 20000,
// width 8, tau_phi 1.0; :49-55: optional argv[1] scales force_z = 4e-6*(1+a)).
#include "microchannel_common.h"
int main(int argc, char **argv)
{
    double force_z = 0.00003;
    if (argc > 1) force_z = 0.000004 + atof(argv[1]) * 0.000004;
    std::cout << force_z << "\n";
    ChannelSetup S = {92, 52, 1500, 250001, 20000, 50, 8, 6, 1.0, force_z};
    return run_channel(S, argc, argv);
}
