// TEST INFRASTRUCTURE -- recording stand-ins for the handful of VTK 7.1 classes the reference's include/VtkOutputter.hpp
// uses (VTK is absent from this image).  They keep what the reference's code hands to VTK -- points, point
// scalars and their name, cells with type and node ids, the file name -- and vtkXMLUnstructuredGridWriter::Write() dumps
// that record to the file in a flat binary form the tests read (NOT VTK's XML: the XML bytes are VTK's, the CONTENT is the
// reference's).  Record layout, little-endian: int64 npoints, ncells, nconn, name_len; double xyz[3*npoints];
// double scalars[npoints]; int64 types[ncells]; int64 offsets[ncells] (end of each cell in conn); int64 conn[nconn];
// char name[name_len] (the active scalars).
#pragma once
#include <cstdio>
#include <memory>
#include <string>
#include <vector>

typedef long long vtkIdType;
enum VTKCellType { VTK_TRIANGLE = 5, VTK_QUAD = 9 };  // vtkCellType.h
#define VTK_MAJOR_VERSION 7

template <class T>
class vtkSmartPointer {
  std::shared_ptr<T> p_;
 public:
  vtkSmartPointer() {}
  static vtkSmartPointer New() { vtkSmartPointer s; s.p_ = std::make_shared<T>(); return s; }
  T* operator->() const { return p_.get(); }
  operator T*() const { return p_.get(); }
};

class vtkPoints {
 public:
  std::vector<double> xyz;
  void InsertPoint(vtkIdType id, double x, double y, double z) {
    if ((size_t)(3 * (id + 1)) > xyz.size()) xyz.resize((size_t)(3 * (id + 1)), 0.0);
    xyz[3 * id] = x; xyz[3 * id + 1] = y; xyz[3 * id + 2] = z;
  }
};
class vtkCellArray {};
class vtkDoubleArray {
 public:
  std::string name;
  std::vector<double> v;
  void SetName(const char* n) { name = n; }
  void SetNumberOfValues(vtkIdType n) { v.assign((size_t)n, 0.0); }
  void SetValue(vtkIdType i, double x) { v.at((size_t)i) = x; }
};
class vtkPointData {
 public:
  vtkDoubleArray* scalars = nullptr;
  std::string active;
  void SetScalars(vtkDoubleArray* a) { scalars = a; }
  void SetActiveScalars(const char* n) { active = n; }
};
class vtkUnstructuredGrid {
 public:
  std::vector<vtkIdType> types, offsets, conn;
  vtkPoints* points = nullptr;
  vtkPointData pd;
  void Allocate(vtkIdType) {}
  vtkIdType InsertNextCell(int type, vtkIdType n, vtkIdType* ids) {
    types.push_back(type);
    conn.insert(conn.end(), ids, ids + n);
    offsets.push_back((vtkIdType)conn.size());
    return (vtkIdType)types.size() - 1;
  }
  void SetPoints(vtkPoints* p) { points = p; }
  vtkPointData* GetPointData() { return &pd; }
};
class vtkXMLUnstructuredGridWriter {
 public:
  std::string file;
  vtkUnstructuredGrid* grid = nullptr;
  const char* GetDefaultFileExtension() { return "vtu"; }  // vtkXMLUnstructuredGridWriter::GetDefaultFileExtension
  void SetFileName(const char* f) { file = f; }
  void SetInputData(vtkUnstructuredGrid* g) { grid = g; }
  int Write() {
    FILE* fh = std::fopen(file.c_str(), "wb");
    if (!fh || !grid || !grid->points || !grid->pd.scalars) return 0;
    const long long hdr[4] = {(long long)(grid->points->xyz.size() / 3), (long long)grid->types.size(), (long long)grid->conn.size(),
                              (long long)grid->pd.active.size()};
    std::fwrite(hdr, sizeof(long long), 4, fh);
    std::fwrite(grid->points->xyz.data(), sizeof(double), grid->points->xyz.size(), fh);
    std::fwrite(grid->pd.scalars->v.data(), sizeof(double), grid->pd.scalars->v.size(), fh);
    std::fwrite(grid->types.data(), sizeof(vtkIdType), grid->types.size(), fh);
    std::fwrite(grid->offsets.data(), sizeof(vtkIdType), grid->offsets.size(), fh);
    std::fwrite(grid->conn.data(), sizeof(vtkIdType), grid->conn.size(), fh);
    std::fwrite(grid->pd.active.data(), 1, grid->pd.active.size(), fh);
    std::fclose(fh);
    return 1;
  }
};
