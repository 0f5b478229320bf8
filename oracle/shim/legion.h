/* oracle/shim/legion.h — TEST INFRASTRUCTURE, not product code.
 *
 * A miniature stand-in for the Legion runtime headers (the reference's deps/legion
 * submodule is empty and Legion is not installed anywhere in this image).  It exists
 * for one purpose: to let the reference's own, UNMODIFIED sources
 *     /root/reference/src/ops/embedding.cc   (CPU task bodies + AVX2 lookup)
 *     /root/reference/src/ops/embedding.cu   (GPU task bodies + embed_forward/backward)
 * and the reference's own, UNMODIFIED headers (include/model.h, config.h, tensor.h,
 * accessor.h, ...) compile in place, so that the parity oracle is the reference's
 * arithmetic itself.  It is placed on the include path AHEAD of nothing else named
 * legion.h (there is none); every other header comes from /root/reference/include.
 *
 * Only the surface those two translation units and the headers they include touch is
 * provided.  Regions are plain host/device pointers plus a rectangle; there is no
 * scheduling, no mapping, no copies: "tasks" are called directly by oracle/ref_glue.
 */
#ifndef FFEMB_ORACLE_LEGION_SHIM_H
#define FFEMB_ORACLE_LEGION_SHIM_H

#include <cassert>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <map>
#include <set>
#include <sstream>
#include <string>
#include <vector>

#ifndef __CUDACC__
#ifndef __host__
#define __host__
#endif
#ifndef __device__
#define __device__
#endif
#endif

namespace Legion {

typedef long long coord_t;
typedef unsigned int FieldID;
typedef unsigned long MappingTagID;
typedef unsigned int TaskID;
typedef unsigned int MapperID;
typedef unsigned int ReductionOpID;
typedef unsigned int ProjectionID;
typedef unsigned int TraceID;
typedef unsigned int VariantID;
typedef unsigned int LayoutConstraintID;
typedef int AddressSpace;

enum PrivilegeMode { NO_ACCESS = 0, READ_ONLY = 1, READ_WRITE = 7, WRITE_ONLY = 2,
                     WRITE_DISCARD = 2, REDUCE = 8 };
enum CoherenceProperty { EXCLUSIVE = 0, ATOMIC = 1, SIMULTANEOUS = 2, RELAXED = 3 };

/* ---- geometry -------------------------------------------------------------------- */
template <int N, typename T = coord_t>
struct Point {
  T x[N];
  __host__ __device__ Point() { for (int i = 0; i < N; i++) x[i] = 0; }
  __host__ __device__ explicit Point(T v) { for (int i = 0; i < N; i++) x[i] = v; }
  __host__ __device__ Point(T a, T b) { static_assert(N == 2, "2-d ctor"); x[0] = a; x[1] = b; }
  __host__ __device__ T &operator[](int i) { return x[i]; }
  __host__ __device__ const T &operator[](int i) const { return x[i]; }
  bool operator==(const Point &o) const {
    for (int i = 0; i < N; i++) if (x[i] != o.x[i]) return false;
    return true;
  }
  bool operator<(const Point &o) const {
    for (int i = 0; i < N; i++) { if (x[i] < o.x[i]) return true; if (x[i] > o.x[i]) return false; }
    return false;
  }
};

class Domain;

template <int N, typename T = coord_t>
struct Rect {
  Point<N, T> lo, hi;
  Rect() {}
  Rect(const Point<N, T> &l, const Point<N, T> &h) : lo(l), hi(h) {}
  Rect(const Domain &d);  /* defined after Domain */
  size_t volume() const {
    size_t v = 1;
    for (int i = 0; i < N; i++) { if (hi[i] < lo[i]) return 0; v *= (size_t)(hi[i] - lo[i] + 1); }
    return v;
  }
  bool operator==(const Rect &o) const { return lo == o.lo && hi == o.hi; }
  bool operator!=(const Rect &o) const { return !(*this == o); }
  bool contains(const Rect &o) const {
    for (int i = 0; i < N; i++) if (o.lo[i] < lo[i] || o.hi[i] > hi[i]) return false;
    return true;
  }
};

class Domain {
public:
  enum { MAX_RECT_DIM = 8 };
  int dim;
  coord_t rect_data[2 * MAX_RECT_DIM];
  Domain() : dim(0) { std::memset(rect_data, 0, sizeof(rect_data)); }
  template <int N, typename T>
  Domain(const Rect<N, T> &r) : dim(N) {
    std::memset(rect_data, 0, sizeof(rect_data));
    for (int i = 0; i < N; i++) { rect_data[i] = r.lo[i]; rect_data[N + i] = r.hi[i]; }
  }
  int get_dim() const { return dim; }
  size_t get_volume() const {
    size_t v = 1;
    for (int i = 0; i < dim; i++) v *= (size_t)(rect_data[dim + i] - rect_data[i] + 1);
    return v;
  }
  bool operator==(const Domain &o) const {
    if (dim != o.dim) return false;
    for (int i = 0; i < 2 * dim; i++) if (rect_data[i] != o.rect_data[i]) return false;
    return true;
  }
  bool operator!=(const Domain &o) const { return !(*this == o); }
  bool operator<(const Domain &o) const {
    if (dim != o.dim) return dim < o.dim;
    for (int i = 0; i < 2 * dim; i++) {
      if (rect_data[i] < o.rect_data[i]) return true;
      if (rect_data[i] > o.rect_data[i]) return false;
    }
    return false;
  }
  struct DomainLoHi {
    const coord_t *p;
    coord_t operator[](int i) const { return p[i]; }
  };
  DomainLoHi lo() const { DomainLoHi r = {rect_data}; return r; }
  DomainLoHi hi() const { DomainLoHi r = {rect_data + dim}; return r; }
};

template <int N, typename T>
Rect<N, T>::Rect(const Domain &d) {
  assert(d.dim == N);
  for (int i = 0; i < N; i++) { lo[i] = (T)d.rect_data[i]; hi[i] = (T)d.rect_data[N + i]; }
}

class DomainPoint {
public:
  int dim;
  coord_t point_data[Domain::MAX_RECT_DIM];
  DomainPoint() : dim(0) { std::memset(point_data, 0, sizeof(point_data)); }
  template <int N, typename T>
  DomainPoint(const Point<N, T> &p) : dim(N) {
    std::memset(point_data, 0, sizeof(point_data));
    for (int i = 0; i < N; i++) point_data[i] = p[i];
  }
  coord_t operator[](int i) const { return point_data[i]; }
  bool operator<(const DomainPoint &o) const {
    if (dim != o.dim) return dim < o.dim;
    for (int i = 0; i < dim; i++) {
      if (point_data[i] < o.point_data[i]) return true;
      if (point_data[i] > o.point_data[i]) return false;
    }
    return false;
  }
};

template <int N, typename T = coord_t>
class PointInRectIterator {
  Rect<N, T> r;
  Point<N, T> p;
  bool valid;
public:
  PointInRectIterator(const Rect<N, T> &rect) : r(rect), p(rect.lo), valid(rect.volume() > 0) {}
  bool operator()() const { return valid; }
  Point<N, T> operator*() const { return p; }
  PointInRectIterator &operator++() {
    for (int i = 0; i < N; i++) {
      if (p[i] < r.hi[i]) { p[i]++; return *this; }
      p[i] = r.lo[i];
    }
    valid = false;
    return *this;
  }
  PointInRectIterator operator++(int) { PointInRectIterator t = *this; ++(*this); return t; }
};

template <int M, int N, typename T = coord_t>
struct Transform { T m[M][N]; T *operator[](int i) { return m[i]; } };

/* ---- handles ----------------------------------------------------------------------- */
class IndexSpace {
public:
  Domain dom;
  unsigned id;
  IndexSpace() : id(0) {}
  explicit IndexSpace(const Domain &d) : dom(d), id(1) {}
  bool operator==(const IndexSpace &o) const { return id == o.id && dom == o.dom; }
  bool operator!=(const IndexSpace &o) const { return !(*this == o); }
  bool operator<(const IndexSpace &o) const { return dom < o.dom; }
};
template <int N, typename T = coord_t>
class IndexSpaceT : public IndexSpace {
public:
  IndexSpaceT() {}
  explicit IndexSpaceT(const IndexSpace &s) : IndexSpace(s) {}
  IndexSpaceT(const Rect<N, T> &r) : IndexSpace(Domain(r)) {}
};
class IndexPartition {
public:
  Domain color_space;
  bool operator==(const IndexPartition &o) const { return color_space == o.color_space; }
};
class FieldSpace { public: unsigned id; FieldSpace() : id(0) {} };
class LogicalRegion {
public:
  IndexSpace is;
  unsigned id;
  LogicalRegion() : id(0) {}
  explicit LogicalRegion(const IndexSpace &s) : is(s), id(1) {}
  IndexSpace get_index_space() const { return is; }
  bool operator==(const LogicalRegion &o) const { return id == o.id && is == o.is; }
  bool operator!=(const LogicalRegion &o) const { return !(*this == o); }
  bool operator<(const LogicalRegion &o) const { return is < o.is; }
  static const LogicalRegion NO_REGION;
};
class LogicalPartition {
public:
  IndexPartition ip;
  IndexPartition get_index_partition() const { return ip; }
  bool operator==(const LogicalPartition &o) const { return ip == o.ip; }
  static const LogicalPartition NO_PART;
};

class Memory {
public:
  enum Kind { NO_MEMKIND, GLOBAL_MEM, SYSTEM_MEM, REGDMA_MEM, SOCKET_MEM, Z_COPY_MEM,
              GPU_FB_MEM, DISK_MEM, HDF_MEM, FILE_MEM, LEVEL3_CACHE, LEVEL2_CACHE, LEVEL1_CACHE };
  unsigned long long id;
  Memory() : id(0) {}
  Kind kind() const { return SYSTEM_MEM; }
  static const Memory NO_MEMORY;
};
class Processor {
public:
  enum Kind { NO_KIND, TOC_PROC, LOC_PROC, UTIL_PROC, IO_PROC, PROC_GROUP, PROC_SET,
              OMP_PROC, PY_PROC };
  unsigned long long id;
  Processor() : id(0) {}
  Kind kind() const { return LOC_PROC; }
};

/* A "physical region" here is a borrowed dense pointer plus the rectangle it covers. */
class PhysicalRegion {
public:
  void *base;
  Domain dom;
  PhysicalRegion() : base(NULL) {}
  PhysicalRegion(void *b, const Domain &d) : base(b), dom(d) {}
  void wait_until_valid() {}
  bool is_mapped() const { return base != NULL; }
};

class RegionRequirement {
public:
  LogicalRegion region;
  LogicalRegion parent;
  LogicalPartition partition;
  PrivilegeMode privilege;
  CoherenceProperty prop;
  ProjectionID projection;
  MappingTagID tag;
  std::set<FieldID> privilege_fields;
  RegionRequirement() : privilege(NO_ACCESS), prop(EXCLUSIVE), projection(0), tag(0) {}
  RegionRequirement(const LogicalRegion &r, PrivilegeMode p, CoherenceProperty c,
                    const LogicalRegion &par, MappingTagID t = 0)
      : region(r), parent(par), privilege(p), prop(c), projection(0), tag(t) {}
  RegionRequirement(const LogicalPartition &lp, ProjectionID proj, PrivilegeMode p,
                    CoherenceProperty c, const LogicalRegion &par, MappingTagID t = 0)
      : parent(par), partition(lp), privilege(p), prop(c), projection(proj), tag(t) {}
  RegionRequirement &add_field(FieldID f) { privilege_fields.insert(f); return *this; }
};

class Task {
public:
  std::vector<RegionRequirement> regions;
  void *args;
  size_t arglen;
  void *local_args;
  size_t local_arglen;
  DomainPoint index_point;
  bool is_index_space;
  MappingTagID tag;
  Task() : args(NULL), arglen(0), local_args(NULL), local_arglen(0), is_index_space(false), tag(0) {}
};

struct ContextStub {};
typedef ContextStub *Context;

class TaskArgument {
public:
  const void *ptr; size_t len;
  TaskArgument() : ptr(NULL), len(0) {}
  TaskArgument(const void *p, size_t l) : ptr(p), len(l) {}
};
/* Like Legion's, the map COPIES the bytes at set_point (the reference's launchers pass the address of a
 * loop-local OpMeta*, src/ops/embedding.cu:281-284). */
class ArgumentMap {
public:
  std::map<DomainPoint, std::vector<char> > bytes;
  template <int N, typename T>
  void set_point(const Point<N, T> &p, const TaskArgument &a) { set_point(DomainPoint(p), a); }
  void set_point(const DomainPoint &p, const TaskArgument &a) {
    const char *b = (const char *)a.ptr;
    bytes[p] = a.len ? std::vector<char>(b, b + a.len) : std::vector<char>();
  }
  size_t size() const { return bytes.size(); }
};
class Predicate {
public:
  static const Predicate TRUE_PRED;
  static const Predicate FALSE_PRED;
};
class Future {
public:
  template <typename T> T get_result() const { return T(); }
  void get_void_result() const {}
  static Future from_value(...) { return Future(); }
};
class FutureMap {
public:
  void wait_all_results() {}
  template <typename T, typename PT> T get_result(const PT &) { return T(); }
  template <typename PT> Future get_future(const PT &) { return Future(); }
};
class TaskLauncher {
public:
  std::vector<RegionRequirement> region_requirements;
  std::vector<Future> futures;
  TaskLauncher() {}
  TaskLauncher(TaskID, TaskArgument, Predicate = Predicate(), MapperID = 0, MappingTagID = 0) {}
  void add_region_requirement(const RegionRequirement &r) { region_requirements.push_back(r); }
  void add_field(unsigned idx, FieldID f) { region_requirements[idx].add_field(f); }
  void add_future(const Future &f) { futures.push_back(f); }
};
/* An index launch keeps what it was built from, so that a test can look at it (Runtime's launch hook
 * below): task id, launch domain, the by-value argument, the per-point arguments, mapper id and tag. */
class IndexLauncher {
public:
  std::vector<RegionRequirement> region_requirements;
  std::vector<Future> futures;
  TaskID task_id;
  Domain launch_domain;
  TaskArgument global_arg;
  ArgumentMap argument_map;
  bool must_parallelism;
  MapperID map_id;
  MappingTagID tag;
  IndexLauncher() : task_id(0), must_parallelism(false), map_id(0), tag(0) {}
  IndexLauncher(TaskID tid, const IndexSpace &space, TaskArgument arg, const ArgumentMap &map,
                Predicate = Predicate(), bool must = false, MapperID mid = 0, MappingTagID t = 0)
      : task_id(tid), launch_domain(space.dom), global_arg(arg), argument_map(map), must_parallelism(must),
        map_id(mid), tag(t) {}
  IndexLauncher(TaskID tid, const Domain &dom, TaskArgument arg, const ArgumentMap &map,
                Predicate = Predicate(), bool must = false, MapperID mid = 0, MappingTagID t = 0)
      : task_id(tid), launch_domain(dom), global_arg(arg), argument_map(map), must_parallelism(must),
        map_id(mid), tag(t) {}
  void add_region_requirement(const RegionRequirement &r) { region_requirements.push_back(r); }
  void add_field(unsigned idx, FieldID f) { region_requirements[idx].add_field(f); }
  void add_future(const Future &f) { futures.push_back(f); }
};
class InlineLauncher {
public:
  InlineLauncher() {}
  InlineLauncher(const RegionRequirement &) {}
};

/* Launches go nowhere (there is no scheduler here) unless a test installs a hook that looks at them. */
class Runtime {
public:
  typedef void (*IndexLaunchHook)(void *user, const IndexLauncher &launcher);
  IndexLaunchHook index_launch_hook;
  void *index_launch_user;
  Runtime() : index_launch_hook(NULL), index_launch_user(NULL) {}
  Domain get_index_space_domain(Context, const IndexSpace &s) { return s.dom; }
  Domain get_index_space_domain(const IndexSpace &s) { return s.dom; }
  Domain get_index_partition_color_space(Context, const IndexPartition &p) { return p.color_space; }
  FutureMap execute_index_space(Context, const IndexLauncher &l) {
    if (index_launch_hook) index_launch_hook(index_launch_user, l);
    return FutureMap();
  }
  Future execute_task(Context, const TaskLauncher &) { return Future(); }
  void begin_trace(Context, TraceID) {}
  void end_trace(Context, TraceID) {}
  void issue_execution_fence(Context) {}
};

typedef Runtime HighLevelRuntime;
}  // namespace Legion

/* ---- Realm bits the reference headers name ----------------------------------------- */
namespace Realm {
template <typename FT, int N, typename T>
struct AffineAccessor {};
class RegionInstance {
public:
  unsigned long long id;
  RegionInstance() : id(0) {}
};
}

namespace Legion {
/* FieldAccessor: .ptr(rect) hands back the dense base pointer of the region. */
template <PrivilegeMode MODE, typename FT, int N, typename T = coord_t,
          typename A = Realm::AffineAccessor<FT, N, T>, bool CB = false>
class FieldAccessor {
  FT *base;
  Domain dom;
public:
  /* the reference asserts acc.accessor.is_dense_arbitrary(rect); shim regions always are */
  struct DenseProbe {
    template <typename R> bool is_dense_arbitrary(const R &) const { return true; }
  } accessor;
  FieldAccessor() : base(NULL) {}
  FieldAccessor(const PhysicalRegion &r, FieldID) : base((FT *)r.base), dom(r.dom) {}
  FT *ptr(const Rect<N, T> &r) const {
    Rect<N, T> mine(dom);
    assert(mine.contains(r));
    /* dense, dim-0-fastest */
    size_t off = 0, stride = 1;
    for (int i = 0; i < N; i++) { off += (size_t)(r.lo[i] - mine.lo[i]) * stride;
                                  stride *= (size_t)(mine.hi[i] - mine.lo[i] + 1); }
    return base + off;
  }
  FT *ptr(const Point<N, T> &p) const { return ptr(Rect<N, T>(p, p)); }
};
}  // namespace Legion

namespace LegionRuntime {
namespace Logger {
class Category {
public:
  Category(const char *) {}
  void print(const char *, ...) {}
  void debug(const char *, ...) {}
  void info(const char *, ...) {}
  void warning(const char *, ...) {}
  void error(const char *, ...) {}
  void spew(const char *, ...) {}
};
}  // namespace Logger
}  // namespace LegionRuntime

/* Legion's dimension dispatch macro; the oracle builds with MAX_TENSOR_DIM=4 */
#ifndef LEGION_FOREACH_N
#define LEGION_FOREACH_N(FFM) FFM(1) FFM(2) FFM(3) FFM(4)
#endif

#endif /* FFEMB_ORACLE_LEGION_SHIM_H */
