// host_common.h -- shared by the drop-in host programs (hosts/*.cpp).
// Each program keeps its reference counterpart's constants, initial condition,
// stdout lines and output files, and replaces the body of the time loop by calls
// into liblbm_b200.so (include/lbm_b200.h).
#ifndef LBM_HOST_COMMON_H
#define LBM_HOST_COMMON_H
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "lbm_b200.h"

#define LBM_CHECK(ctx, call)                                                                  \
    do {                                                                                      \
        lbm_status _s = (call);                                                               \
        if (_s != LBM_OK) {                                                                   \
            std::fprintf(stderr, "%s failed (%d): %s\n", #call, (int)_s, lbm_last_error(ctx)); \
            std::exit(1);                                                                     \
        }                                                                                     \
    } while (0)

inline int device_from_env()
{
    const char *e = std::getenv("LBM_DEVICE");
    return e ? std::atoi(e) : 0;
}

// "name" + zero padding to `width` digits + counter, as the reference builds its file names
inline std::string padded(const std::string &stem, int width, int counter)
{
    std::stringstream num;
    num << counter;
    std::stringstream out;
    out << stem << std::string(width - (int)num.str().size(), '0') << counter;
    return out.str();
}
#endif
