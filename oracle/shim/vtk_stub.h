/* Stand-in for the seven VTK-5 headers Microchannel3D includes (VTK is not in
 * the image).  The writers/readers are never called by the oracle driver; the
 * classes only have to let the reference's templates parse.  VTK 5's own
 * vtkIOStream.h put std::cout / std::endl into the global namespace and the
 * reference relies on that (FullChannel/main.cpp:145) as well as on a global
 * isnan (solver.hpp:52), so this header provides them too.  Test infrastructure
 * only. */
#ifndef LBM_ORACLE_SHIM_VTK_STUB_H
#define LBM_ORACLE_SHIM_VTK_STUB_H
#include <cmath>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
using std::cout;
using std::endl;
using std::isnan;

typedef long long vtkIdType;

template <class T>
class vtkSmartPointer {
public:
    vtkSmartPointer() : p_(new T) {}
    static vtkSmartPointer New() { return vtkSmartPointer(); }
    T *operator->() const { return p_; }
    operator T *() const { return p_; }
private:
    T *p_;  // leaked on purpose: never instantiated at run time
};

class vtkIdList {};
class vtkPoints {
public:
    template <class... A> void InsertNextPoint(A...) {}
    vtkIdType GetNumberOfPoints() { return 0; }
};
class vtkDataArray {
public:
    vtkIdType GetNumberOfTuples() { return 0; }
    double *GetTuple(vtkIdType) { return 0; }
};
class vtkDoubleArray : public vtkDataArray {
public:
    void SetNumberOfComponents(int) {}
    void SetName(const char *) {}
    void InsertNextTupleValue(const double *) {}
};
class vtkPointData {
public:
    void AddArray(vtkDataArray *) {}
    vtkDataArray *GetScalars(const char *) { return 0; }
    vtkDataArray *GetVectors(const char *) { return 0; }
};
class vtkStructuredGrid {
public:
    static vtkStructuredGrid *New() { return new vtkStructuredGrid; }
    void SetDimensions(int, int, int) {}
    void SetPoints(vtkPoints *) {}
    vtkPointData *GetPointData() { return &pd_; }
    void ShallowCopy(vtkStructuredGrid *) {}
private:
    vtkPointData pd_;
};
class vtkXMLStructuredGridWriter {
public:
    void SetFileName(const char *) {}
    void SetInput(vtkStructuredGrid *) {}
    int Write() { return 1; }
};
class vtkXMLStructuredGridReader {
public:
    void SetFileName(const char *) {}
    void Update() {}
    vtkStructuredGrid *GetOutput() { return 0; }
};
#endif
