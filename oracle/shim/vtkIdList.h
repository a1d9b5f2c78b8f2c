/* forwards to the single stub; see vtk_stub.h */
#include "vtk_stub.h"
