// Kokkos_Core.hpp — a host-only stand-in for the part of the Kokkos programming model that the reference's
// P3 headers (components/eamxx/src/physics/p3/impl/*.hpp, physics/share/*.hpp) and the EKAT headers they include
// (externals/ekat/src/ekat/) use.  TEST INFRASTRUCTURE ONLY: it exists so that the reference's OWN, UNMODIFIED
// source files compile with g++ into oracle/_ref/libp3ref*.so (recipe: oracle/Makefile.ref), which pins the oracle
// (DESIGN.md §4).  Kokkos itself is an empty sub-module in the reference tree (externals/ekat/extern/kokkos).
//
// Nothing here is derived from Kokkos source: it is a fresh, minimal implementation of the documented interface
// (View with layouts / memory traits / static extents, subview, deep_copy, create_mirror_view, TeamPolicy with
// a one-thread team per league member, nested TeamVectorRange loops, single, the Max/Min/Sum reducers, Array,
// pair).  Semantics follow the Kokkos OpenMP host back end with team size 1 — the configuration EKAT chooses on
// CPUs (ekat_kokkos_utils.hpp:224-231): league members run as iterations of an OpenMP loop, everything nested in
// a team runs sequentially in index order.
#ifndef P3REF_KOKKOS_CORE_SHIM_HPP
#define P3REF_KOKKOS_CORE_SHIM_HPP

#include <algorithm>
#include <array>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <new>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#define KOKKOS_INLINE_FUNCTION inline
#define KOKKOS_FORCEINLINE_FUNCTION inline __attribute__((always_inline))
#define KOKKOS_FUNCTION
#define KOKKOS_LAMBDA [=]
#define KOKKOS_CLASS_LAMBDA [=, *this]
#define KOKKOS_COMPILER_GNU 1
#define KOKKOS_ENABLE_OPENMP 1
#define KOKKOS_VERSION 40000

namespace Kokkos {

// ------------------------------------------------------------------------------------------------ spaces
struct HostSpace {
  using memory_space = HostSpace;
  using size_type = std::size_t;
  static constexpr const char* name() { return "Host"; }
};
struct ScratchMemorySpaceHost {
  using memory_space = ScratchMemorySpaceHost;
  char* base = nullptr;
  std::size_t bytes = 0;
  mutable std::size_t used = 0;
  void* get(std::size_t n) const {
    n = (n + 7) & ~std::size_t(7);
    assert(used + n <= bytes);
    void* p = base + used;
    used += n;
    return p;
  }
};
struct OpenMP {
  using execution_space = OpenMP;
  using memory_space = HostSpace;
  using scratch_memory_space = ScratchMemorySpaceHost;
  using size_type = std::size_t;
  int concurrency() const {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
  }
  static int impl_thread_pool_size() { return OpenMP().concurrency(); }
  void fence() const {}
  static const char* name() { return "OpenMP(shim)"; }
};
using Serial = OpenMP;   // one host back end is enough here; with one thread it IS the serial back end
using DefaultExecutionSpace = OpenMP;
using DefaultHostExecutionSpace = OpenMP;

template <class ExecSpace, class MemSpace>
struct Device {
  using execution_space = ExecSpace;
  using memory_space = MemSpace;
  using device_type = Device;
};

struct LayoutRight {};
struct LayoutLeft {};
struct LayoutStride {};

enum MemoryTraitsFlags : unsigned { Unmanaged = 0x01, RandomAccess = 0x02, Atomic = 0x04, Restrict = 0x08, Aligned = 0x10 };
template <unsigned T>
struct MemoryTraits {
  enum : bool {
    is_unmanaged = (T & unsigned(Unmanaged)) != 0,
    is_random_access = (T & unsigned(RandomAccess)) != 0,
    is_atomic = (T & unsigned(Atomic)) != 0,
    is_restrict = (T & unsigned(Restrict)) != 0,
    is_aligned = (T & unsigned(Aligned)) != 0
  };
  static constexpr unsigned value = T;
};
using MemoryManaged = MemoryTraits<0>;
using MemoryUnmanaged = MemoryTraits<Unmanaged>;
using MemoryRandomAccess = MemoryTraits<Unmanaged | RandomAccess>;

inline void fence() {}
inline void fence(const std::string&) {}
inline void memory_fence() {}
[[noreturn]] inline void abort(const char* msg) {
  std::fprintf(stderr, "Kokkos::abort: %s\n", msg);
  std::abort();
}
inline bool is_initialized() { return true; }

template <class T>
inline bool atomic_compare_exchange_strong(T* dest, T expected, T desired) {
  return __atomic_compare_exchange_n(dest, &expected, desired, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
}

// ------------------------------------------------------------------------------------------------ small utilities
template <class T, std::size_t N>
struct Array {
  T m_data[N ? N : 1];
  using value_type = T;
  using reference = T&;
  using const_reference = const T&;
  static constexpr std::size_t size() { return N; }
  T& operator[](std::size_t i) { return m_data[i]; }
  const T& operator[](std::size_t i) const { return m_data[i]; }
  T* data() { return m_data; }
  const T* data() const { return m_data; }
  T* begin() { return m_data; }
  T* end() { return m_data + N; }
  const T* begin() const { return m_data; }
  const T* end() const { return m_data + N; }
};

template <class A, class B>
struct pair {
  A first;
  B second;
  pair() = default;
  pair(const A& a, const B& b) : first(a), second(b) {}
  template <class C, class D>
  pair(const pair<C, D>& o) : first(o.first), second(o.second) {}
  template <class C, class D>
  pair(const std::pair<C, D>& o) : first(o.first), second(o.second) {}
};
template <class A, class B>
pair<A, B> make_pair(A a, B b) { return pair<A, B>(a, b); }

struct ALL_t {};
constexpr ALL_t ALL{};

template <class T>
struct reduction_identity {
  static constexpr T sum() { return T(0); }
  static constexpr T prod() { return T(1); }
  static constexpr T max() { return std::numeric_limits<T>::lowest(); }
  static constexpr T min() { return std::numeric_limits<T>::max(); }
  static constexpr T lor() { return T(0); }
  static constexpr T land() { return T(1); }
};

struct ViewAllocateWithoutInitializing {
  std::string label;
  explicit ViewAllocateWithoutInitializing(const std::string& l) : label(l) {}
};

// ------------------------------------------------------------------------------------------------ View
namespace Impl {

// DataType analysis: "T**[3][4]" -> value T, two run-time extents followed by static extents 3, 4.
template <class T>
struct PtrDepth { static constexpr int value = 0; using base = T; };
template <class T>
struct PtrDepth<T*> { static constexpr int value = 1 + PtrDepth<T>::value; using base = typename PtrDepth<T>::base; };

template <class DT, int I, int R>
struct StaticExtents {
  static void fill(std::size_t* e) {
    e[I] = std::extent<DT, I>::value;
    StaticExtents<DT, I + 1, R>::fill(e);
  }
};
template <class DT, int R>
struct StaticExtents<DT, R, R> { static void fill(std::size_t*) {} };

template <class DT>
struct DataAnalysis {
  using noext = typename std::remove_all_extents<DT>::type;                     // e.g. "const double" or "double**"
  static constexpr int rank_static = int(std::rank<DT>::value);
  static constexpr int rank_dynamic = PtrDepth<noext>::value;
  static constexpr int rank = rank_static + rank_dynamic;
  using value_type = typename PtrDepth<noext>::base;                           // keeps const
  using non_const_value_type = typename std::remove_const<value_type>::type;
  static void static_extents(std::size_t* e) { StaticExtents<DT, 0, rank_static>::fill(e + rank_dynamic); }
};

// re-target a data type to another value type, e.g. const double[5][4] -> double[5][4]
template <class DT, class NewV>
struct Rebind { using type = NewV; };
template <class DT, class NewV>
struct Rebind<DT*, NewV> { using type = typename Rebind<DT, NewV>::type*; };
template <class DT, class NewV, std::size_t N>
struct Rebind<DT[N], NewV> { using type = typename Rebind<DT, NewV>::type[N]; };

// property classification
template <class P> struct is_layout : std::false_type {};
template <> struct is_layout<LayoutRight> : std::true_type {};
template <> struct is_layout<LayoutLeft> : std::true_type {};
template <> struct is_layout<LayoutStride> : std::true_type {};
template <class P> struct is_memtraits : std::false_type {};
template <unsigned T> struct is_memtraits<MemoryTraits<T>> : std::true_type {};

template <class Default, template <class> class Pred, class... Ps>
struct FindProp { using type = Default; };
template <class Default, template <class> class Pred, class P, class... Ps>
struct FindProp<Default, Pred, P, Ps...> {
  using type = typename std::conditional<Pred<P>::value, P, typename FindProp<Default, Pred, Ps...>::type>::type;
};

template <class P, class = void> struct is_space : std::false_type {};
template <class P>
struct is_space<P, typename std::enable_if<!is_layout<P>::value && !is_memtraits<P>::value>::type> : std::true_type {};

template <class S, class = void>
struct SpaceToDevice { using type = Device<DefaultExecutionSpace, S>; };   // a bare memory space
template <class S>
struct SpaceToDevice<S, typename std::enable_if<!std::is_same<typename S::execution_space, void>::value>::type> {
  using type = Device<typename S::execution_space, typename S::memory_space>;
};

struct AllocRecord {
  void* ptr = nullptr;
  std::string label;
  ~AllocRecord() { std::free(ptr); }
};

struct SubAll {};  // markers for subview arguments
}  // namespace Impl

template <class DataType, class... Props>
struct ViewTraits {
  using data_type = DataType;
  using analysis = Impl::DataAnalysis<DataType>;
  using value_type = typename analysis::value_type;
  using const_value_type = typename std::add_const<value_type>::type;
  using non_const_value_type = typename analysis::non_const_value_type;
  using scalar_array_type = DataType;
  using const_data_type = typename Impl::Rebind<DataType, const_value_type>::type;
  using non_const_data_type = typename Impl::Rebind<DataType, non_const_value_type>::type;
  using array_layout = typename Impl::FindProp<LayoutRight, Impl::is_layout, Props...>::type;
  using memory_traits = typename Impl::FindProp<MemoryManaged, Impl::is_memtraits, Props...>::type;
  using specified_space = typename Impl::FindProp<Device<DefaultExecutionSpace, HostSpace>, Impl::is_space, Props...>::type;
  using device_type = typename Impl::SpaceToDevice<specified_space>::type;
  using execution_space = typename device_type::execution_space;
  using memory_space = typename device_type::memory_space;
  using size_type = std::size_t;
  static constexpr int rank = analysis::rank;
  static constexpr int rank_dynamic = analysis::rank_dynamic;
  static constexpr bool is_managed = !memory_traits::is_unmanaged;
};

template <class DataType, class... Props>
class View {
 public:
  using traits = ViewTraits<DataType, Props...>;
  using data_type = DataType;
  using value_type = typename traits::value_type;
  using const_value_type = typename traits::const_value_type;
  using non_const_value_type = typename traits::non_const_value_type;
  using reference_type = value_type&;
  using pointer_type = value_type*;
  using array_layout = typename traits::array_layout;
  using memory_traits = typename traits::memory_traits;
  using device_type = typename traits::device_type;
  using execution_space = typename traits::execution_space;
  using memory_space = typename traits::memory_space;
  using size_type = std::size_t;
  using non_const_type = View<typename traits::non_const_data_type, array_layout, device_type, memory_traits>;
  using const_type = View<typename traits::const_data_type, array_layout, device_type, memory_traits>;
  using HostMirror = View<typename traits::non_const_data_type, array_layout, Device<DefaultHostExecutionSpace, HostSpace>>;
  using host_mirror_type = HostMirror;
  static constexpr int rank = traits::rank;
  static constexpr int Rank = traits::rank;
  static constexpr int rank_dynamic = traits::rank_dynamic;
  static constexpr bool is_layout_right = std::is_same<array_layout, LayoutRight>::value;
  static constexpr bool is_layout_left = std::is_same<array_layout, LayoutLeft>::value;
  static constexpr int NR = rank ? rank : 1;

 private:
  template <class, class...> friend class View;
  pointer_type m_ptr = nullptr;
  std::size_t m_ext[NR] = {};
  std::size_t m_str[NR] = {};
  std::shared_ptr<Impl::AllocRecord> m_rec;

  void set_extents(const std::size_t* dyn, int ndyn) {
    (void)ndyn;
    for (int i = 0; i < rank_dynamic; ++i) m_ext[i] = dyn[i];
    traits::analysis::static_extents(m_ext);
    compute_strides();
  }
  void compute_strides() {
    if (rank == 0) return;
    if (is_layout_left) {
      std::size_t s = 1;
      for (int i = 0; i < rank; ++i) { m_str[i] = s; s *= m_ext[i]; }
    } else {
      std::size_t s = 1;
      for (int i = rank - 1; i >= 0; --i) { m_str[i] = s; s *= m_ext[i]; }
    }
  }
  void allocate(const std::string& label, bool init) {
    static_assert(!std::is_const<value_type>::value || true, "");
    const std::size_t n = size();
    auto rec = std::make_shared<Impl::AllocRecord>();
    rec->label = label;
    const std::size_t bytes = std::max<std::size_t>(n * sizeof(value_type), 64);
    rec->ptr = std::aligned_alloc(64, (bytes + 63) / 64 * 64);
    if (!rec->ptr) throw std::bad_alloc();
    using NC = non_const_value_type;
    NC* p = static_cast<NC*>(rec->ptr);
    if (init) {
      if (std::is_trivial<NC>::value) std::memset(rec->ptr, 0, n * sizeof(NC));
      else for (std::size_t i = 0; i < n; ++i) new (p + i) NC();
    } else if (!std::is_trivially_default_constructible<NC>::value) {
      for (std::size_t i = 0; i < n; ++i) new (p + i) NC;
    }
    m_ptr = p;
    m_rec = std::move(rec);
  }
  template <std::size_t... Is, class... I>
  std::size_t offset(std::index_sequence<Is...>, I... idx) const {
    const std::size_t ii[] = {static_cast<std::size_t>(idx)...};
    std::size_t o = 0;
    for (int d = 0; d < rank; ++d) {
      assert(ii[d] < m_ext[d]);
      o += ii[d] * ((is_layout_right && d == rank - 1) ? std::size_t(1) : m_str[d]);
    }
    return o;
  }

 public:
  View() = default;
  View(const View&) = default;
  View(View&&) = default;
  View& operator=(const View&) = default;
  View& operator=(View&&) = default;

  // allocating constructors
  template <class... Dims, class = typename std::enable_if<(std::is_integral<Dims>::value && ...)>::type>
  explicit View(const std::string& label, Dims... dims) {
    const std::size_t d[] = {static_cast<std::size_t>(dims)..., 0};
    set_extents(d, int(sizeof...(Dims)));
    allocate(label, true);
  }
  template <class... Dims, class = typename std::enable_if<(std::is_integral<Dims>::value && ...)>::type>
  explicit View(const char* label, Dims... dims) : View(std::string(label), dims...) {}
  template <class... Dims, class = typename std::enable_if<(std::is_integral<Dims>::value && ...)>::type>
  explicit View(const ViewAllocateWithoutInitializing& a, Dims... dims) {
    const std::size_t d[] = {static_cast<std::size_t>(dims)..., 0};
    set_extents(d, int(sizeof...(Dims)));
    allocate(a.label, false);
  }
  // wrapping constructors (unmanaged memory)
  template <class... Dims, class = typename std::enable_if<(std::is_integral<Dims>::value && ...)>::type>
  View(pointer_type p, Dims... dims) : m_ptr(p) {
    const std::size_t d[] = {static_cast<std::size_t>(dims)..., 0};
    set_extents(d, int(sizeof...(Dims)));
  }
  // scratch-memory constructor
  template <class... Dims, class = typename std::enable_if<(std::is_integral<Dims>::value && ...)>::type>
  View(const ScratchMemorySpaceHost& s, Dims... dims) {
    const std::size_t d[] = {static_cast<std::size_t>(dims)..., 0};
    set_extents(d, int(sizeof...(Dims)));
    m_ptr = static_cast<pointer_type>(s.get(size() * sizeof(value_type)));
  }
  // converting constructor: same rank, value types equal up to added const, any layout / traits whose strides fit
  template <class DT2, class... P2,
            class = typename std::enable_if<
                View<DT2, P2...>::rank == rank &&
                std::is_same<typename View<DT2, P2...>::non_const_value_type, non_const_value_type>::value &&
                (std::is_const<value_type>::value || !std::is_const<typename View<DT2, P2...>::value_type>::value) &&
                !std::is_same<View<DT2, P2...>, View>::value>::type>
  View(const View<DT2, P2...>& o) : m_ptr(o.m_ptr) {
    for (int i = 0; i < rank; ++i) { m_ext[i] = o.m_ext[i]; m_str[i] = o.m_str[i]; }
    if (rank > 0) {
      std::size_t chk[NR];
      for (int i = 0; i < rank_dynamic; ++i) chk[i] = m_ext[i];
      traits::analysis::static_extents(chk);
      for (int i = rank_dynamic; i < rank; ++i) assert(chk[i] == m_ext[i] && "static extent mismatch in View conversion");
      if (is_layout_right) assert(m_ext[rank - 1] <= 1 || m_str[rank - 1] == 1);
      if (is_layout_left) assert(m_ext[0] <= 1 || m_str[0] == 1);
    }
    if (traits::is_managed) m_rec = o.m_rec;
  }
  template <class DT2, class... P2,
            class = typename std::enable_if<
                View<DT2, P2...>::rank == rank &&
                std::is_same<typename View<DT2, P2...>::non_const_value_type, non_const_value_type>::value &&
                (std::is_const<value_type>::value || !std::is_const<typename View<DT2, P2...>::value_type>::value) &&
                !std::is_same<View<DT2, P2...>, View>::value>::type>
  View& operator=(const View<DT2, P2...>& o) {
    *this = View(o);
    return *this;
  }

  // raw construction used by subview()
  static View impl_from_parts(pointer_type p, const std::size_t* ext, const std::size_t* str) {
    View v;
    v.m_ptr = p;
    for (int i = 0; i < rank; ++i) { v.m_ext[i] = ext[i]; v.m_str[i] = str[i]; }
    return v;
  }

  template <class... I>
  typename std::enable_if<(sizeof...(I) == std::size_t(rank)) && (rank > 0), reference_type>::type
  operator()(I... idx) const {
    return m_ptr[offset(std::make_index_sequence<sizeof...(I)>(), idx...)];
  }
  template <int R = rank>
  typename std::enable_if<R == 0, reference_type>::type operator()() const { return *m_ptr; }
  template <class I, int R = rank>
  typename std::enable_if<R == 1, reference_type>::type operator[](I i) const { return (*this)(i); }

  template <class I> std::size_t extent(I i) const { return int(i) < rank ? m_ext[int(i)] : 1; }
  template <class I> int extent_int(I i) const { return int(extent(i)); }
  template <class I> std::size_t stride(I i) const { return int(i) < rank ? m_str[int(i)] : 0; }
  std::size_t stride_0() const { return stride(0); }
  std::size_t stride_1() const { return stride(1); }
  std::size_t stride_2() const { return stride(2); }
  std::size_t size() const {
    std::size_t n = 1;
    for (int i = 0; i < rank; ++i) n *= m_ext[i];
    return n;
  }
  std::size_t span() const {
    if (rank == 0) return 1;
    std::size_t s = 0;
    for (int i = 0; i < rank; ++i) {
      if (m_ext[i] == 0) return 0;
      s += (m_ext[i] - 1) * m_str[i];
    }
    return s + 1;
  }
  bool span_is_contiguous() const { return span() == size(); }
  pointer_type data() const { return m_ptr; }
  bool is_allocated() const { return m_ptr != nullptr; }
  std::string label() const { return m_rec ? m_rec->label : std::string(); }
  int use_count() const { return m_rec ? int(m_rec.use_count()) : 0; }
  static std::size_t shmem_size(std::size_t n0 = 1, std::size_t n1 = 1, std::size_t n2 = 1) {
    return ((n0 * n1 * n2 * sizeof(value_type)) + 7) / 8 * 8 + 8;
  }

  // the tiny part of the mapping interface that ekat_subview_utils.hpp reaches into
  struct ImplMap {
    const View* v;
    template <class... I>
    reference_type reference(I... idx) const { return v->m_ptr[v->offset_unchecked(idx...)]; }
  };
  ImplMap impl_map() const { return ImplMap{this}; }
  template <class... I>
  std::size_t offset_unchecked(I... idx) const {
    const std::size_t ii[] = {static_cast<std::size_t>(idx)..., 0};
    std::size_t o = 0;
    for (int d = 0; d < rank && d < int(sizeof...(I)); ++d) o += ii[d] * m_str[d];
    return o;
  }
};

template <class V> struct is_view : std::false_type {};
template <class D, class... P> struct is_view<View<D, P...>> : std::true_type {};
template <class D, class... P> struct is_view<const View<D, P...>> : std::true_type {};

// ------------------------------------------------------------------------------------------------ subview
namespace Impl {
template <class A> struct KeepsDim : std::true_type {};                  // ALL or a range keeps the dimension
template <> struct KeepsDim<int> : std::false_type {};
template <> struct KeepsDim<unsigned> : std::false_type {};
template <> struct KeepsDim<long> : std::false_type {};
template <> struct KeepsDim<unsigned long> : std::false_type {};
template <> struct KeepsDim<long long> : std::false_type {};
template <> struct KeepsDim<unsigned long long> : std::false_type {};
template <> struct KeepsDim<short> : std::false_type {};

template <class V, int N> struct AddPtrs { using type = typename AddPtrs<V, N - 1>::type*; };
template <class V> struct AddPtrs<V, 0> { using type = V; };

inline void sub_arg(std::size_t ext, ALL_t, std::size_t& b, std::size_t& n) { b = 0; n = ext; }
template <class A, class B>
inline void sub_arg(std::size_t, const Kokkos::pair<A, B>& p, std::size_t& b, std::size_t& n) { b = p.first; n = p.second - p.first; }
template <class A, class B>
inline void sub_arg(std::size_t, const std::pair<A, B>& p, std::size_t& b, std::size_t& n) { b = p.first; n = p.second - p.first; }
template <class I>
inline typename std::enable_if<std::is_integral<I>::value>::type
sub_arg(std::size_t, I i, std::size_t& b, std::size_t& n) { b = std::size_t(i); n = 0; }
}  // namespace Impl

// Result: an unmanaged LayoutStride view of the kept dimensions (converts to LayoutRight/Left views when the
// strides allow it; the conversion asserts).
template <class D, class... P, class... Args>
auto subview(const View<D, P...>& v, Args... args) {
  using Src = View<D, P...>;
  static_assert(sizeof...(Args) == std::size_t(Src::rank), "subview: one argument per dimension");
  constexpr int newrank = (0 + ... + (Impl::KeepsDim<Args>::value ? 1 : 0));
  using NewData = typename Impl::AddPtrs<typename Src::value_type, newrank>::type;
  using Dst = View<NewData, LayoutStride, typename Src::device_type, MemoryTraits<Unmanaged>>;
  std::size_t b[Src::NR], n[Src::NR];
  const bool keep[] = {Impl::KeepsDim<Args>::value..., false};
  int d = 0;
  auto step = [&](auto a) { Impl::sub_arg(v.extent(d), a, b[d], n[d]); ++d; };
  (step(args), ...);
  std::size_t off = 0, ext[Dst::NR] = {}, str[Dst::NR] = {};
  int k = 0;
  for (int i = 0; i < Src::rank; ++i) {
    off += b[i] * v.stride(i);
    if (keep[i]) { ext[k] = n[i]; str[k] = v.stride(i); ++k; }
  }
  return Dst::impl_from_parts(v.data() + off, ext, str);
}

// ------------------------------------------------------------------------------------------------ mirrors, deep_copy
template <class D, class... P>
typename View<D, P...>::HostMirror create_mirror(const View<D, P...>& v) {
  using M = typename View<D, P...>::HostMirror;
  return [&]() {
    if constexpr (M::rank_dynamic == 0) return M("mirror");
    else if constexpr (M::rank_dynamic == 1) return M("mirror", v.extent(0));
    else if constexpr (M::rank_dynamic == 2) return M("mirror", v.extent(0), v.extent(1));
    else if constexpr (M::rank_dynamic == 3) return M("mirror", v.extent(0), v.extent(1), v.extent(2));
    else if constexpr (M::rank_dynamic == 4) return M("mirror", v.extent(0), v.extent(1), v.extent(2), v.extent(3));
    else return M("mirror", v.extent(0), v.extent(1), v.extent(2), v.extent(3), v.extent(4));
  }();
}
// host memory is the only memory: a mirror VIEW of a non-const view is the view itself (Kokkos semantics when the
// memory spaces coincide); a const view gets a fresh copy target
template <class D, class... P>
typename View<D, P...>::HostMirror create_mirror_view(const View<D, P...>& v) {
  using V = View<D, P...>;
  if constexpr (!std::is_const<typename V::value_type>::value && !std::is_same<typename V::array_layout, LayoutStride>::value) {
    return typename V::HostMirror(v);
  } else {
    return create_mirror(v);
  }
}

namespace Impl {
template <class Dst, class Src>
void copy_rec(const Dst& d, const Src& s, std::size_t* idx, int dim) {
  constexpr int R = Dst::rank;
  if (dim == R) {
    std::size_t od = 0, os = 0;
    for (int i = 0; i < R; ++i) { od += idx[i] * d.stride(i); os += idx[i] * s.stride(i); }
    d.data()[od] = s.data()[os];
    return;
  }
  for (std::size_t i = 0; i < d.extent(dim); ++i) { idx[dim] = i; copy_rec(d, s, idx, dim + 1); }
}
template <class Dst, class T>
void fill_rec(const Dst& d, const T& val, std::size_t* idx, int dim) {
  constexpr int R = Dst::rank;
  if (dim == R) {
    std::size_t od = 0;
    for (int i = 0; i < R; ++i) od += idx[i] * d.stride(i);
    d.data()[od] = val;
    return;
  }
  for (std::size_t i = 0; i < d.extent(dim); ++i) { idx[dim] = i; fill_rec(d, val, idx, dim + 1); }
}
}  // namespace Impl

template <class D1, class... P1, class D2, class... P2>
void deep_copy(const View<D1, P1...>& dst, const View<D2, P2...>& src) {
  static_assert(View<D1, P1...>::rank == View<D2, P2...>::rank, "deep_copy: rank mismatch");
  if ((const void*)dst.data() == (const void*)src.data()) return;
  for (int i = 0; i < View<D1, P1...>::rank; ++i) assert(dst.extent(i) == src.extent(i));
  std::size_t idx[View<D1, P1...>::NR] = {};
  if (dst.size() == 0) return;
  Impl::copy_rec(dst, src, idx, 0);
}
template <class D1, class... P1>
void deep_copy(const View<D1, P1...>& dst, const typename View<D1, P1...>::non_const_value_type& val) {
  std::size_t idx[View<D1, P1...>::NR] = {};
  if (dst.size() == 0) return;
  Impl::fill_rec(dst, val, idx, 0);
}
template <class E, class A, class B>
void deep_copy(const E&, const A& a, const B& b) { deep_copy(a, b); }

// ------------------------------------------------------------------------------------------------ policies
struct PerTeamValue { std::size_t value; };
struct PerThreadValue { std::size_t value; };
inline PerTeamValue PerTeam(std::size_t v) { return PerTeamValue{v}; }
inline PerThreadValue PerThread(std::size_t v) { return PerThreadValue{v}; }
struct AUTO_t {};
constexpr AUTO_t AUTO{};

class HostTeamMember {
  int m_league_rank, m_league_size;
  ScratchMemorySpaceHost m_scratch;

 public:
  HostTeamMember(int lr, int ls, char* scratch, std::size_t bytes) : m_league_rank(lr), m_league_size(ls) {
    m_scratch.base = scratch;
    m_scratch.bytes = bytes;
  }
  int league_rank() const { return m_league_rank; }
  int league_size() const { return m_league_size; }
  int team_rank() const { return 0; }
  int team_size() const { return 1; }
  void team_barrier() const {}
  const ScratchMemorySpaceHost& team_scratch(int) const { return m_scratch; }
  const ScratchMemorySpaceHost& team_shmem() const { return m_scratch; }
  const ScratchMemorySpaceHost& thread_scratch(int) const { return m_scratch; }
  template <class T> void team_broadcast(T&, int) const {}
};

template <class... Props>
class TeamPolicy {
  int m_league = 0, m_team = 1, m_vector = 1;
  std::size_t m_scratch[2] = {0, 0};

 public:
  using member_type = HostTeamMember;
  using execution_space = DefaultExecutionSpace;
  TeamPolicy() = default;
  TeamPolicy(int league, int team, int vector = 1) : m_league(league), m_team(team), m_vector(vector) {}
  TeamPolicy(int league, AUTO_t, int vector = 1) : m_league(league), m_vector(vector) {}
  template <class E, class = typename std::enable_if<!std::is_integral<E>::value>::type>
  TeamPolicy(const E&, int league, int team, int vector = 1) : m_league(league), m_team(team), m_vector(vector) {}
  int league_size() const { return m_league; }
  int team_size() const { return 1; }          // host back end: EKAT asks for 1 thread per team on CPUs anyway
  int requested_team_size() const { return m_team; }
  TeamPolicy& set_scratch_size(int level, PerTeamValue v) { m_scratch[level ? 1 : 0] = v.value; return *this; }
  TeamPolicy& set_scratch_size(int level, PerTeamValue v, PerThreadValue t) { m_scratch[level ? 1 : 0] = v.value + t.value; return *this; }
  TeamPolicy& set_scratch_size(int level, PerThreadValue t) { m_scratch[level ? 1 : 0] = t.value; return *this; }
  std::size_t scratch_bytes() const { return m_scratch[0] + m_scratch[1]; }
  template <class F> int team_size_max(const F&, ...) const { return 1; }
  template <class F> int team_size_recommended(const F&, ...) const { return 1; }
};

template <class... Props>
class RangePolicy {
  std::int64_t m_b = 0, m_e = 0;

 public:
  using execution_space = DefaultExecutionSpace;
  using member_type = std::int64_t;
  RangePolicy() = default;
  RangePolicy(std::int64_t b, std::int64_t e) : m_b(b), m_e(e) {}
  template <class E, class = typename std::enable_if<!std::is_integral<E>::value>::type>
  RangePolicy(const E&, std::int64_t b, std::int64_t e) : m_b(b), m_e(e) {}
  std::int64_t begin() const { return m_b; }
  std::int64_t end() const { return m_e; }
};

struct NestedRange { int begin, end; };
template <class M> NestedRange TeamVectorRange(const M&, int n) { return NestedRange{0, n}; }
template <class M> NestedRange TeamVectorRange(const M&, int b, int e) { return NestedRange{b, e}; }
template <class M> NestedRange TeamThreadRange(const M&, int n) { return NestedRange{0, n}; }
template <class M> NestedRange TeamThreadRange(const M&, int b, int e) { return NestedRange{b, e}; }
template <class M> NestedRange ThreadVectorRange(const M&, int n) { return NestedRange{0, n}; }
template <class M> NestedRange ThreadVectorRange(const M&, int b, int e) { return NestedRange{b, e}; }

struct SingleToken {};
template <class M> SingleToken PerTeam(const M&) { return SingleToken{}; }
template <class M> SingleToken PerThread(const M&) { return SingleToken{}; }
template <class F> void single(SingleToken, const F& f) { f(); }
template <class F, class T> void single(SingleToken, const F& f, T& val) { f(val); }

// reducers
template <class T, class Space = HostSpace>
struct Max {
  using value_type = T;
  T& ref;
  explicit Max(T& r) : ref(r) {}
  void init(T& v) const { v = reduction_identity<T>::max(); }
  void join(T& d, const T& s) const { if (s > d) d = s; }
  T& reference() const { return ref; }
};
template <class T, class Space = HostSpace>
struct Min {
  using value_type = T;
  T& ref;
  explicit Min(T& r) : ref(r) {}
  void init(T& v) const { v = reduction_identity<T>::min(); }
  void join(T& d, const T& s) const { if (s < d) d = s; }
  T& reference() const { return ref; }
};
template <class T, class Space = HostSpace>
struct Sum {
  using value_type = T;
  T& ref;
  explicit Sum(T& r) : ref(r) {}
  void init(T& v) const { v = reduction_identity<T>::sum(); }
  void join(T& d, const T& s) const { d += s; }
  T& reference() const { return ref; }
};
namespace Impl {
template <class R, class = void> struct is_reducer : std::false_type {};
template <class R> struct is_reducer<R, std::void_t<typename R::value_type, decltype(std::declval<const R&>().reference())>> : std::true_type {};
}  // namespace Impl

// nested (inside a team): sequential, in index order
template <class F> void parallel_for(const NestedRange& r, const F& f) {
  for (int i = r.begin; i < r.end; ++i) f(i);
}
template <class F, class T>
typename std::enable_if<!Impl::is_reducer<T>::value>::type
parallel_reduce(const NestedRange& r, const F& f, T& result) {
  T acc = reduction_identity<T>::sum();
  for (int i = r.begin; i < r.end; ++i) f(i, acc);
  result = acc;
}
template <class F, class R>
typename std::enable_if<Impl::is_reducer<R>::value>::type
parallel_reduce(const NestedRange& r, const F& f, const R& red) {
  typename R::value_type acc;
  red.init(acc);
  for (int i = r.begin; i < r.end; ++i) f(i, acc);
  red.reference() = acc;
}

// top level
template <class... P, class F>
void parallel_for(const std::string&, const TeamPolicy<P...>& pol, const F& f) {
  const int n = pol.league_size();
  const std::size_t sb = pol.scratch_bytes() + 64;
#ifdef _OPENMP
#pragma omp parallel
#endif
  {
    std::vector<char> scratch(sb);
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
    for (int i = 0; i < n; ++i) {
      HostTeamMember m(i, n, scratch.data(), sb);
      f(m);
    }
  }
}
template <class... P, class F>
void parallel_for(const TeamPolicy<P...>& pol, const F& f) { parallel_for(std::string(), pol, f); }
template <class... P, class F>
void parallel_for(const std::string&, const RangePolicy<P...>& pol, const F& f) {
  for (std::int64_t i = pol.begin(); i < pol.end(); ++i) f(i);
}
template <class... P, class F>
void parallel_for(const RangePolicy<P...>& pol, const F& f) { parallel_for(std::string(), pol, f); }
template <class I, class F>
typename std::enable_if<std::is_integral<I>::value>::type
parallel_for(const std::string&, I n, const F& f) { for (I i = 0; i < n; ++i) f(i); }
template <class I, class F>
typename std::enable_if<std::is_integral<I>::value>::type
parallel_for(I n, const F& f) { for (I i = 0; i < n; ++i) f(i); }

template <class... P, class F, class T>
typename std::enable_if<!Impl::is_reducer<T>::value>::type
parallel_reduce(const std::string&, const RangePolicy<P...>& pol, const F& f, T& result) {
  T acc = reduction_identity<T>::sum();
  for (std::int64_t i = pol.begin(); i < pol.end(); ++i) f(i, acc);
  result = acc;
}
template <class... P, class F, class T>
typename std::enable_if<!Impl::is_reducer<T>::value>::type
parallel_reduce(const RangePolicy<P...>& pol, const F& f, T& result) { parallel_reduce(std::string(), pol, f, result); }
template <class I, class F, class T>
typename std::enable_if<std::is_integral<I>::value && !Impl::is_reducer<T>::value>::type
parallel_reduce(const std::string&, I n, const F& f, T& result) {
  T acc = reduction_identity<T>::sum();
  for (I i = 0; i < n; ++i) f(i, acc);
  result = acc;
}
template <class I, class F, class T>
typename std::enable_if<std::is_integral<I>::value && !Impl::is_reducer<T>::value>::type
parallel_reduce(I n, const F& f, T& result) { parallel_reduce(std::string(), n, f, result); }

inline void initialize(int&, char**) {}
inline void initialize() {}
inline void finalize() {}

}  // namespace Kokkos

#endif  // P3REF_KOKKOS_CORE_SHIM_HPP
