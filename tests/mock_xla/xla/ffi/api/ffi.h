// TEST INFRASTRUCTURE ONLY — a minimal stand-in for jaxlib's xla/ffi/api/ffi.h (absent from this image), with just
// the names trex_b200/csrc/xla_ffi_shim.cc uses, so that `g++ -fsyntax-only` can type-check the shim's handler
// bodies (tests/test_abi_and_host.py).  It binds nothing and proves nothing about the real XLA FFI semantics.
#pragma once
#include <cstddef>
#include <cstdint>
#include <optional>
#include <vector>
namespace xla { namespace ffi {
enum DataType { F32, F64, S32 };
template <DataType> struct NativeOf;
template <> struct NativeOf<F32> { using type = float; };
template <> struct NativeOf<F64> { using type = double; };
template <> struct NativeOf<S32> { using type = int32_t; };
struct Span {
  std::vector<int64_t> v;
  size_t size() const { return v.size(); }
  int64_t operator[](size_t i) const { return v[i]; }
};
template <DataType T> struct Buffer {
  using N = typename NativeOf<T>::type;
  Span dims; N* ptr = nullptr;
  const Span& dimensions() const { return dims; }
  N* typed_data() const { return ptr; }
};
template <DataType T> struct ResultBuffer {
  Buffer<T> b;
  Buffer<T>* operator->() { return &b; }
};
struct ScratchAllocator { std::optional<void*> Allocate(size_t) { return std::nullopt; } };
struct Error {
  static Error Success() { return {}; }
  static Error InvalidArgument(const char*) { return {}; }
  static Error Internal(const char*) { return {}; }
};
template <typename S> struct PlatformStream {};
struct Binding {
  template <typename T> Binding& Ctx() { return *this; }
  template <typename T> Binding& Arg() { return *this; }
  template <typename T> Binding& Ret() { return *this; }
  template <typename T> Binding& Attr(const char*) { return *this; }
};
struct Ffi { static Binding Bind() { return {}; } };
}}  // namespace xla::ffi
#define XLA_FFI_DEFINE_HANDLER_SYMBOL(name, impl, binding) \
  extern "C" void* name() { (void)(binding); return (void*)&impl; }
