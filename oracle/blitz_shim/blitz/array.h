// TEST INFRASTRUCTURE -- a minimal stand-in for <blitz/array.h> (blitz++ 1.0.1 is absent from this image).
//
// Purpose: let oracle/make_ref.py compile the reference's OWN sw2d::computeRHS / computeTimeStep / buildSpongeCoeff
// (reference src/sw2d/main.cpp:251-484, 517-553 and src/sw2d-simple/main.cpp:179-356), together with the reference's own
// Types.hpp, BlitzHelpers.hpp, DGContext2D.hpp, BCtypes.hpp and Constants.hpp, VERBATIM from where they lie under
// /root/reference, into oracle/_ref/libsw2d_ref.so.  Only what those lines use is provided, with blitz++'s semantics:
//   * Array<T,N>, N = 1, 2: row-major (C) storage by default; the copy constructor ALIASES the storage (reference
//     counting), operator= copies element values;
//   * whole-array expressions are evaluated element by element, in C operator order, one statement at a time;
//   * sum(expr, thirdIndex) is a sequential left-to-right sum starting from T(0); max(expr) a sequential scan with '>';
//   * index placeholders (firstIndex ...) evaluate to the loop index; A(ii,kk) re-maps a rank-2 array into a tensor
//     expression; the extent of a reduced index comes from the arrays that carry it.
// No blitz++ source was consulted or copied: this restates the documented behaviour the call sites rely on.
#ifndef BDG_BLITZ_SHIM_ARRAY_H
#define BDG_BLITZ_SHIM_ARRAY_H
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <iostream>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

namespace blitz {

template <int N>
struct IndexPlaceholder;
template <typename E>
struct ETBase {
  const E &self() const { return static_cast<const E &>(*this); }
};

// shape bookkeeping of an expression: ext[d] = extent of tensor index d, or -1 if no operand carries it
struct Shape {
  int ext[3];
  Shape() { ext[0] = ext[1] = ext[2] = -1; }
  void merge(const Shape &o) {
    for (int d = 0; d < 3; ++d)
      if (o.ext[d] >= 0) {
        if (ext[d] >= 0 && ext[d] != o.ext[d]) throw std::runtime_error("blitz shim: shape mismatch in expression");
        ext[d] = o.ext[d];
      }
  }
};

template <int N>
struct IndexPlaceholder : ETBase<IndexPlaceholder<N>> {
  typedef int value_type;
  int eval(const int *idx) const { return idx[N]; }
  Shape shape() const { return Shape(); }
};
typedef IndexPlaceholder<0> firstIndex;
typedef IndexPlaceholder<1> secondIndex;
typedef IndexPlaceholder<2> thirdIndex;
namespace tensor {
static const firstIndex i = firstIndex();
static const secondIndex j = secondIndex();
static const thirdIndex k = thirdIndex();
}  // namespace tensor

struct Range {
  int lo, hi;
  Range(int a, int b) : lo(a), hi(b) {}
  static Range all() { return Range(0, -1); }
};
template <int N>
struct GeneralArrayStorage {
  bool colMajor;
  GeneralArrayStorage() : colMajor(false) {}
};
template <int N>
struct ColumnMajorArray : GeneralArrayStorage<N> {
  ColumnMajorArray() { this->colMajor = true; }
};
typedef ColumnMajorArray<2> FortranArray2;

template <typename T>
struct Constant : ETBase<Constant<T>> {
  typedef T value_type;
  T v;
  explicit Constant(T x) : v(x) {}
  T eval(const int *) const { return v; }
  Shape shape() const { return Shape(); }
};

#define BDG_BINOP(NAME, OP)                                                                          \
  template <typename L, typename R>                                                                  \
  struct NAME : ETBase<NAME<L, R>> {                                                                 \
    typedef decltype(typename L::value_type() OP typename R::value_type()) value_type;               \
    L l;                                                                                             \
    R r;                                                                                             \
    NAME(const L &a, const R &b) : l(a), r(b) {}                                                     \
    value_type eval(const int *idx) const { return l.eval(idx) OP r.eval(idx); }                     \
    Shape shape() const { Shape s = l.shape(); s.merge(r.shape()); return s; }                      \
  };
BDG_BINOP(AddExpr, +)
BDG_BINOP(SubExpr, -)
BDG_BINOP(MulExpr, *)
BDG_BINOP(DivExpr, /)
BDG_BINOP(LtExpr, <)  // comparisons give 0 / 1 per element (zf = (abs(r) < 1.0-1e-10), src/TriangleNodesProvisioner.cpp:637)
BDG_BINOP(GtExpr, >)
BDG_BINOP(ModExpr, %)  // integer arrays only (f1 = (f1 % NumFaces), src/MeshManager.cpp:465)
#undef BDG_BINOP

template <typename E>
struct NegExpr : ETBase<NegExpr<E>> {
  typedef typename E::value_type value_type;
  E e;
  explicit NegExpr(const E &x) : e(x) {}
  value_type eval(const int *idx) const { return -e.eval(idx); }
  Shape shape() const { return e.shape(); }
};
template <typename E>
struct SqrtExpr : ETBase<SqrtExpr<E>> {
  typedef double value_type;
  E e;
  explicit SqrtExpr(const E &x) : e(x) {}
  double eval(const int *idx) const { return std::sqrt(e.eval(idx)); }
  Shape shape() const { return e.shape(); }
};
template <typename E>
struct FloorExpr : ETBase<FloorExpr<E>> {  // floor(expr), element by element (of an integer quotient it is the identity)
  typedef double value_type;
  E e;
  explicit FloorExpr(const E &x) : e(x) {}
  double eval(const int *idx) const { return std::floor((double)e.eval(idx)); }
  Shape shape() const { return e.shape(); }
};
template <typename E>
struct PowExpr : ETBase<PowExpr<E>> {  // pow(expr, scalar), element by element (std::pow on doubles)
  typedef double value_type;
  E e;
  double p;
  PowExpr(const E &x, double q) : e(x), p(q) {}
  double eval(const int *idx) const { return std::pow((double)e.eval(idx), p); }
  Shape shape() const { return e.shape(); }
};
template <typename E>
struct AbsExpr : ETBase<AbsExpr<E>> {
  typedef typename E::value_type value_type;
  E e;
  explicit AbsExpr(const E &x) : e(x) {}
  value_type eval(const int *idx) const { return std::abs(e.eval(idx)); }
  Shape shape() const { return e.shape(); }
};
// sum(expr, IndexPlaceholder<D>): sequential sum over index D starting from 0
template <typename E, int D>
struct SumExpr : ETBase<SumExpr<E, D>> {
  typedef typename E::value_type value_type;
  E e;
  int n;
  explicit SumExpr(const E &x) : e(x), n(x.shape().ext[D]) {
    if (n < 0) throw std::runtime_error("blitz shim: reduced index has no extent");
  }
  value_type eval(const int *idx) const {
    int id[3] = {idx[0], idx[1], idx[2]};
    value_type acc = value_type(0);
    for (int k = 0; k < n; ++k) {
      id[D] = k;
      acc += e.eval(id);
    }
    return acc;
  }
  Shape shape() const { Shape s = e.shape(); s.ext[D] = -1; return s; }
};

// max(expr, IndexPlaceholder<D>): sequential scan over index D with '>' from the lowest value
template <typename E, int D>
struct MaxExpr : ETBase<MaxExpr<E, D>> {
  typedef typename E::value_type value_type;
  E e;
  int n;
  explicit MaxExpr(const E &x) : e(x), n(x.shape().ext[D]) {
    if (n < 0) throw std::runtime_error("blitz shim: reduced index has no extent");
  }
  value_type eval(const int *idx) const {
    int id[3] = {idx[0], idx[1], idx[2]};
    value_type m = std::numeric_limits<value_type>::lowest();
    for (int k = 0; k < n; ++k) {
      id[D] = k;
      const value_type x = e.eval(id);
      if (x > m) m = x;
    }
    return m;
  }
  Shape shape() const { Shape s = e.shape(); s.ext[D] = -1; return s; }
};

// operand holders: arrays are held by reference-counted alias, expressions by value
template <typename T, int N>
class Array;
template <typename T, int N, int P0, int P1>
struct IndexedArray;
template <typename T, int P>
struct IndexedVector;

template <typename E> struct is_et : std::is_base_of<ETBase<E>, E> {};

#define BDG_OPS(NAME, OP)                                                                                                  \
  template <typename L, typename R>                                                                                        \
  typename std::enable_if<is_et<L>::value && is_et<R>::value, NAME<L, R>>::type operator OP(const L &a, const R &b) {      \
    return NAME<L, R>(a, b);                                                                                               \
  }                                                                                                                        \
  template <typename L, typename S>                                                                                        \
  typename std::enable_if<is_et<L>::value && std::is_arithmetic<S>::value, NAME<L, Constant<S>>>::type operator OP(        \
      const L &a, S b) {                                                                                                   \
    return NAME<L, Constant<S>>(a, Constant<S>(b));                                                                        \
  }                                                                                                                        \
  template <typename S, typename R>                                                                                        \
  typename std::enable_if<std::is_arithmetic<S>::value && is_et<R>::value, NAME<Constant<S>, R>>::type operator OP(        \
      S a, const R &b) {                                                                                                   \
    return NAME<Constant<S>, R>(Constant<S>(a), b);                                                                        \
  }
BDG_OPS(AddExpr, +)
BDG_OPS(SubExpr, -)
BDG_OPS(MulExpr, *)
BDG_OPS(DivExpr, /)
BDG_OPS(LtExpr, <)
BDG_OPS(GtExpr, >)
BDG_OPS(ModExpr, %)
#undef BDG_OPS
template <typename E>
typename std::enable_if<is_et<E>::value, NegExpr<E>>::type operator-(const E &e) { return NegExpr<E>(e); }
template <typename E>
typename std::enable_if<is_et<E>::value, SqrtExpr<E>>::type sqrt(const E &e) { return SqrtExpr<E>(e); }
template <typename E>
typename std::enable_if<is_et<E>::value, AbsExpr<E>>::type abs(const E &e) { return AbsExpr<E>(e); }
template <typename E>
typename std::enable_if<is_et<E>::value, FloorExpr<E>>::type floor(const E &e) { return FloorExpr<E>(e); }
template <typename E, typename S>
typename std::enable_if<is_et<E>::value && std::is_arithmetic<S>::value, PowExpr<E>>::type pow(const E &e, S p) { return PowExpr<E>(e, (double)p); }
template <typename E, int D>
typename std::enable_if<is_et<E>::value, SumExpr<E, D>>::type sum(const E &e, IndexPlaceholder<D>) { return SumExpr<E, D>(e); }

template <typename E, int D>
typename std::enable_if<is_et<E>::value, MaxExpr<E, D>>::type max(const E &e, IndexPlaceholder<D>) { return MaxExpr<E, D>(e); }

// full reductions over an expression (rank from its shape), row-major traversal
template <typename E, typename F>
void for_each_index(const E &e, F f) {
  const Shape s = e.shape();
  const int n0 = s.ext[0] < 0 ? 1 : s.ext[0], n1 = s.ext[1] < 0 ? 1 : s.ext[1], n2 = s.ext[2] < 0 ? 1 : s.ext[2];
  int idx[3];
  for (idx[0] = 0; idx[0] < n0; ++idx[0])
    for (idx[1] = 0; idx[1] < n1; ++idx[1])
      for (idx[2] = 0; idx[2] < n2; ++idx[2]) f(e.eval(idx));
}
template <typename E>
typename std::enable_if<is_et<E>::value, typename E::value_type>::type max(const E &e) {
  typedef typename E::value_type T;
  T m = std::numeric_limits<T>::lowest();
  for_each_index(e, [&m](T x) { if (x > m) m = x; });
  return m;
}
template <typename E>
typename std::enable_if<is_et<E>::value, typename E::value_type>::type min(const E &e) {
  typedef typename E::value_type T;
  T m = std::numeric_limits<T>::max();
  for_each_index(e, [&m](T x) { if (x < m) m = x; });
  return m;
}
template <typename E>
typename std::enable_if<is_et<E>::value, typename E::value_type>::type sum(const E &e) {
  typedef typename E::value_type T;
  T acc = T(0);
  for_each_index(e, [&acc](T x) { acc += x; });
  return acc;
}

// `A = v0, v1, v2;` (blitz list initialisation): scalar assignment hands back this proxy; a following comma turns the fill into
// an element list in storage order, otherwise the destructor fills the whole array with the scalar (blitz's
// ListInitializationSwitch).
template <typename T, int N>
struct ListInit {
  Array<T, N> *a;
  T first;
  int pos;
  bool list;
  ListInit(Array<T, N> *arr, T f) : a(arr), first(f), pos(1), list(false) {}
  ListInit(ListInit &&o) : a(o.a), first(o.first), pos(o.pos), list(o.list) { o.a = nullptr; }
  ListInit(const ListInit &) = delete;
  ~ListInit();
  template <typename S>
  typename std::enable_if<std::is_arithmetic<S>::value, ListInit &>::type operator,(S x);
  operator Array<T, N> &();
};

template <typename T, int N>
class Array : public ETBase<Array<T, N>> {
  static_assert(N >= 1 && N <= 3, "blitz shim: rank 1..3");
  template <typename, int> friend class Array;
  std::shared_ptr<std::vector<T>> store_;
  T *data_;
  int ext_[3];
  int stride_[3];
  bool colMajor_;

  void alloc(int n0, int n1, int n2, bool colMajor) {
    ext_[0] = n0; ext_[1] = n1; ext_[2] = n2;
    colMajor_ = colMajor;
    const size_t total = (size_t)std::max(n0, 0) * std::max(n1, 1) * std::max(n2, 1);
    store_ = std::make_shared<std::vector<T>>(total ? total : 1);
    data_ = store_->data();
    if (N == 1) { stride_[0] = 1; stride_[1] = stride_[2] = 0; }
    else if (N == 2) {
      if (colMajor) { stride_[0] = 1; stride_[1] = n0; } else { stride_[0] = n1; stride_[1] = 1; }
      stride_[2] = 0;
    } else { stride_[0] = n1 * n2; stride_[1] = n2; stride_[2] = 1; }
  }

 public:
  typedef T value_type;
  typedef T T_numtype;
  typedef T *iterator;
  typedef const T *const_iterator;

  Array() { alloc(0, N > 1 ? 0 : 1, 1, false); }
  explicit Array(int n0) { alloc(n0, 1, 1, false); }
  Array(int n0, int n1) { alloc(n0, n1, 1, false); }
  Array(int n0, int n1, int n2) { alloc(n0, n1, n2, false); }
  template <int M>
  Array(int n0, int n1, GeneralArrayStorage<M> st) { alloc(n0, n1, 1, st.colMajor); }
  Array(const Array &o) : ETBase<Array<T, N>>(), store_(o.store_), data_(o.data_), colMajor_(o.colMajor_) {  // ALIASES the storage
    for (int d = 0; d < 3; ++d) { ext_[d] = o.ext_[d]; stride_[d] = o.stride_[d]; }
  }
  template <typename E, typename = typename std::enable_if<is_et<E>::value && !std::is_same<E, Array>::value>::type>
  Array(const E &e) {  // construction from an expression allocates and evaluates
    const Shape s = e.shape();
    alloc(s.ext[0], N > 1 ? s.ext[1] : 1, N > 2 ? s.ext[2] : 1, false);
    assign(e, [](T &a, T b) { a = b; });
  }

  void resize(int n0) { alloc(n0, 1, 1, false); }
  void resize(int n0, int n1) { alloc(n0, n1, 1, colMajor_); }
  void reference(const Array &o) {
    store_ = o.store_; data_ = o.data_; colMajor_ = o.colMajor_;
    for (int d = 0; d < 3; ++d) { ext_[d] = o.ext_[d]; stride_[d] = o.stride_[d]; }
  }
  Array copy() const {
    Array r;
    r.alloc(ext_[0], ext_[1], ext_[2], colMajor_);
    std::copy(store_->begin(), store_->end(), r.store_->begin());
    return r;
  }

  int rows() const { return ext_[0]; }
  int cols() const { return ext_[1]; }
  int depth() const { return ext_[2]; }
  int length(int d) const { return ext_[d]; }
  struct Extents {  // x.length() of blitz returns a TinyVector; the reference reads (x.length())(0)
    int e[3];
    int operator()(int d) const { return e[d]; }
    int operator[](int d) const { return e[d]; }
  };
  Extents length() const { return Extents{{ext_[0], ext_[1], ext_[2]}}; }
  explicit Array(const Extents &e) { alloc(e.e[0], N > 1 ? e.e[1] : 1, N > 2 ? e.e[2] : 1, false); }  // real_vector_type p(x.length())
  int extent(int d) const { return ext_[d]; }
  int size() const { return numElements(); }
  int numElements() const { return N == 1 ? ext_[0] : (N == 2 ? ext_[0] * ext_[1] : ext_[0] * ext_[1] * ext_[2]); }
  bool isMajorRank(int d) const { return N == 1 ? d == 0 : (colMajor_ ? d == N - 1 : d == 0); }
  T *data() { return data_; }
  const T *data() const { return data_; }
  T *dataFirst() { return data_; }
  const T *dataFirst() const { return data_; }
  // iteration in storage order (the reference iterates rank-1 arrays and row-major matrices only)
  iterator begin() { return data_; }
  iterator end() { return data_ + numElements(); }
  const_iterator begin() const { return data_; }
  const_iterator end() const { return data_ + numElements(); }

  T &operator()(int i) { return data_[(size_t)i * stride_[0]]; }
  const T &operator()(int i) const { return data_[(size_t)i * stride_[0]]; }
  T &operator()(int i, int j) { return data_[(size_t)i * stride_[0] + (size_t)j * stride_[1]]; }
  const T &operator()(int i, int j) const { return data_[(size_t)i * stride_[0] + (size_t)j * stride_[1]]; }
  T &operator()(int i, int j, int k) { return data_[(size_t)i * stride_[0] + (size_t)j * stride_[1] + (size_t)k * stride_[2]]; }
  const T &operator()(int i, int j, int k) const { return data_[(size_t)i * stride_[0] + (size_t)j * stride_[1] + (size_t)k * stride_[2]]; }
  T &operator[](int i) { return data_[i]; }
  const T &operator[](int i) const { return data_[i]; }
  template <int P0, int P1>
  IndexedArray<T, N, P0, P1> operator()(IndexPlaceholder<P0>, IndexPlaceholder<P1>) const { return IndexedArray<T, N, P0, P1>(*this); }
  // v(ii) / v(jj): a rank-1 array riding on tensor index P
  template <int P>
  IndexedVector<T, P> operator()(IndexPlaceholder<P>) const {
    static_assert(N == 1, "blitz shim: single-placeholder indexing of vectors only");
    return IndexedVector<T, P>(*this);
  }
  // A(Range, Range): rank-2 VIEW of a block of rows / columns; T3(Range::all(), Range::all(), k): matrix view of a rank-3 array
  Array<T, 2> operator()(Range r0, Range r1) const {
    static_assert(N == 2, "blitz shim: (Range, Range) on matrices");
    const int a0 = r0.lo, b0 = r0.hi < 0 ? ext_[0] - 1 : r0.hi, a1 = r1.lo, b1 = r1.hi < 0 ? ext_[1] - 1 : r1.hi;
    Array<T, 2> v;
    v.store_ = store_;
    v.data_ = data_ + (size_t)a0 * stride_[0] + (size_t)a1 * stride_[1];
    v.ext_[0] = b0 - a0 + 1; v.ext_[1] = b1 - a1 + 1; v.ext_[2] = 1;
    v.stride_[0] = stride_[0]; v.stride_[1] = stride_[1]; v.stride_[2] = 0;
    v.colMajor_ = colMajor_;
    return v;
  }
  Array<T, 2> operator()(Range, Range, int k) const {
    static_assert(N == 3, "blitz shim: (Range, Range, int) on rank-3 arrays");
    Array<T, 2> v;
    v.store_ = store_;
    v.data_ = data_ + (size_t)k * stride_[2];
    v.ext_[0] = ext_[0]; v.ext_[1] = ext_[1]; v.ext_[2] = 1;
    v.stride_[0] = stride_[0]; v.stride_[1] = stride_[1]; v.stride_[2] = 0;
    v.colMajor_ = false;
    return v;
  }
  // A(i, Range::all()) / A(Range::all(), j): rank-1 VIEWS of a row / column of a matrix (aliases, like blitz)
  Array<T, 1> operator()(int i, Range) const { return slice1(data_ + (size_t)i * stride_[0], ext_[1], stride_[1]); }
  Array<T, 1> operator()(Range, int j) const { return slice1(data_ + (size_t)j * stride_[1], ext_[0], stride_[0]); }

 private:
  Array<T, 1> slice1(T *first, int n, int stride) const {
    static_assert(N == 2, "blitz shim: slices of matrices only");
    Array<T, 1> v;
    v.store_ = store_;
    v.data_ = first;
    v.ext_[0] = n; v.ext_[1] = 1; v.ext_[2] = 1;
    v.stride_[0] = stride; v.stride_[1] = v.stride_[2] = 0;
    return v;
  }

 public:

  // as an expression operand: rank-1 arrays live on tensor index 0, rank-2 on (0, 1)
  T eval(const int *idx) const { return N == 1 ? (*this)(idx[0]) : (N == 2 ? (*this)(idx[0], idx[1]) : (*this)(idx[0], idx[1], idx[2])); }
  Shape shape() const {
    Shape s;
    for (int d = 0; d < N; ++d) s.ext[d] = ext_[d];
    return s;
  }

  template <typename E, typename F>
  void assign(const E &e, F f) {
    const Shape s = e.shape();
    for (int d = 0; d < N; ++d)
      // an expression LARGER than its destination is evaluated over the destination's index range, as blitz does without
      // BZ_DEBUG (the reference's weak-derivative block assigns an Nc x Nc product to an Nc x Np matrix,
      // src/TriangleNodesProvisioner.cpp:495-501); a smaller one would be read out of bounds and stays an error
      if (s.ext[d] >= 0 && s.ext[d] < ext_[d])
        throw std::runtime_error("blitz shim: shape mismatch in assignment: destination " + std::to_string(ext_[0]) + " x " + std::to_string(ext_[1]) +
                                 ", expression " + std::to_string(s.ext[0]) + " x " + std::to_string(s.ext[1]) + " x " + std::to_string(s.ext[2]));
    int idx[3] = {0, 0, 0};
    if (N == 1) {
      for (idx[0] = 0; idx[0] < ext_[0]; ++idx[0]) f((*this)(idx[0]), (T)e.eval(idx));
    } else if (N == 2) {
      for (idx[0] = 0; idx[0] < ext_[0]; ++idx[0])
        for (idx[1] = 0; idx[1] < ext_[1]; ++idx[1]) f((*this)(idx[0], idx[1]), (T)e.eval(idx));
    } else {
      for (idx[0] = 0; idx[0] < ext_[0]; ++idx[0])
        for (idx[1] = 0; idx[1] < ext_[1]; ++idx[1])
          for (idx[2] = 0; idx[2] < ext_[2]; ++idx[2]) f((*this)(idx[0], idx[1], idx[2]), (T)e.eval(idx));
    }
  }
  Array &operator=(const Array &o) {  // copies VALUES (unlike the copy constructor)
    if (numElements() == 0 && o.numElements() != 0) alloc(o.ext_[0], o.ext_[1], o.ext_[2], o.colMajor_);
    assign(o, [](T &a, T b) { a = b; });
    return *this;
  }
#define BDG_ASSIGN(OP)                                                                                      \
  template <typename E>                                                                                     \
  typename std::enable_if<is_et<E>::value, Array &>::type operator OP(const E &e) {                         \
    assign(e, [](T &a, T b) { a OP b; });                                                                   \
    return *this;                                                                                           \
  }                                                                                                         \
  template <typename S>                                                                                     \
  typename std::enable_if<std::is_arithmetic<S>::value, Array &>::type operator OP(S v) {                   \
    assign(Constant<T>((T)v), [](T &a, T b) { a OP b; });                                                   \
    return *this;                                                                                           \
  }
  template <typename E>
  typename std::enable_if<is_et<E>::value, Array &>::type operator=(const E &e) {
    assign(e, [](T &a, T b) { a = b; });
    return *this;
  }
  template <typename S>
  typename std::enable_if<std::is_arithmetic<S>::value, ListInit<T, N>>::type operator=(S v) { return ListInit<T, N>(this, (T)v); }
  void fill(T v) { assign(Constant<T>(v), [](T &a, T b) { a = b; }); }
  BDG_ASSIGN(+=)
  BDG_ASSIGN(-=)
  BDG_ASSIGN(*=)
  BDG_ASSIGN(/=)
#undef BDG_ASSIGN
};

template <typename T, int N>
ListInit<T, N>::~ListInit() {
  if (a && !list) a->fill(first);
}
template <typename T, int N>
template <typename S>
typename std::enable_if<std::is_arithmetic<S>::value, ListInit<T, N> &>::type ListInit<T, N>::operator,(S x) {
  if (!list) {
    (*a)[0] = first;
    list = true;
  }
  if (pos < a->numElements()) (*a)[pos] = (T)x;  // a list longer than the array (blitz would write past its end) is cut off
  ++pos;
  return *this;
}
template <typename T, int N>
ListInit<T, N>::operator Array<T, N> &() {
  if (!list) a->fill(first);
  list = true;
  return *a;
}

// A(ii,kk): rank-2 array whose dimensions ride on tensor indices P0, P1
template <typename T, int N, int P0, int P1>
struct IndexedArray : ETBase<IndexedArray<T, N, P0, P1>> {
  // (a rank-1 array indexed with two placeholders -- an off-path inline member of the reference's DGContext2D.hpp does
  // that -- rides on the first placeholder only)
  typedef T value_type;
  Array<T, N> a;  // alias
  explicit IndexedArray(const Array<T, N> &x) : a(x) {}
  T eval(const int *idx) const { return N == 1 ? a(idx[P0]) : a(idx[P0], idx[P1]); }
  Shape shape() const {
    Shape s;
    s.ext[P0] = a.rows();
    if (N > 1) s.ext[P1] = a.cols();
    return s;
  }
};

template <typename T, int P>
struct IndexedVector : ETBase<IndexedVector<T, P>> {
  typedef T value_type;
  Array<T, 1> a;  // alias
  explicit IndexedVector(const Array<T, 1> &x) : a(x) {}
  T eval(const int *idx) const { return a(idx[P]); }
  Shape shape() const {
    Shape s;
    s.ext[P] = a.rows();
    return s;
  }
};

template <typename T, int N>
std::ostream &operator<<(std::ostream &os, const Array<T, N> &a) {
  os << "(" << a.rows();
  if (N > 1) os << " x " << a.cols();
  os << ")\n[ ";
  for (int i = 0; i < a.numElements(); ++i) os << a.data()[i] << " ";
  return os << "]\n";
}

}  // namespace blitz
#endif
