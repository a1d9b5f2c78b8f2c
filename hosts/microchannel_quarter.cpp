// Drop-in for Microchannel3D/QuarterChannel/main.cpp (:18-31: 27x27x52, 1001 steps, output every 100, width 6).
#include "microchannel_common.h"
int main(int argc, char **argv)
{
    ChannelSetup S = {27, 27, 52, 1001, 100, 20, 6, 5, 2.0, 0.00003};
    return run_channel(S, argc, argv);
}
