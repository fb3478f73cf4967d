// TEST INFRASTRUCTURE ONLY (oracle/): stand-ins for the CGAL typedefs of the
// reference's memsurfer/src/Types.hpp:46-59 so that the reference's own
// TriMesh.cpp + TriMesh_kde.cpp compile UNMODIFIED, from where they lie under
// /root/reference, without CGAL (which is absent in this image).
// Nothing here is reference source; the types are dummies that the hot path
// (need_normals / need_pointareas / kde) never touches.
#pragma once
#include <utility>
#include <cstddef>
#include <iostream>
#include <sstream>
#include <cfloat>
#include <cmath>
#include <functional>
#include <numeric>
#include <algorithm>
#include <stdexcept>
#include <vector>
#include <string>

struct Vector3 { double v[3]; Vector3(double a=0,double b=0,double c=0){v[0]=a;v[1]=b;v[2]=c;} double operator[](int i) const {return v[i];} };
struct Point3  { double v[3]; Point3(double a=0,double b=0,double c=0){v[0]=a;v[1]=b;v[2]=c;} double operator[](int i) const {return v[i];} };
inline Vector3 operator-(const Point3 &a, const Point3 &b) { return Vector3(a[0]-b[0], a[1]-b[1], a[2]-b[2]); }
struct Point2  { double v[2]; Point2(double a=0,double b=0){v[0]=a;v[1]=b;} double operator[](int i) const {return v[i];} };
struct Polyhedron {};
namespace CGAL { inline double scalar_product(const Vector3 &a, const Vector3 &b) { return a[0]*b[0]+a[1]*b[1]+a[2]*b[2]; } }
typedef std::pair<Point2, size_t> Point_with_idx;
typedef std::pair<Point3, Vector3> Point_with_normal;
