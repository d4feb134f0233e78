// pbr/Generator.hpp -- minimal std::generator stand-in.
// The reference's Scene::iterateAllNodes() returns std::generator<Node const&> (Scene.hpp:93,
// <generator>, C++23) which g++ 13 does not ship; this is a small coroutine generator with the same
// use (range-for over a lazily produced sequence of references).
#pragma once
#include <coroutine>
#include <exception>
#include <iterator>
#include <memory>
#include <type_traits>
#include <utility>

namespace pbr {
template <typename Ref> class generator {
  static_assert(std::is_reference_v<Ref>, "pbr::generator yields references");
  using Ptr = std::add_pointer_t<std::remove_reference_t<Ref>>;

public:
  struct promise_type {
    Ptr current {};
    std::exception_ptr error {};
    auto get_return_object() -> generator { return generator {std::coroutine_handle<promise_type>::from_promise(*this)}; }
    auto initial_suspend() noexcept -> std::suspend_always { return {}; }
    auto final_suspend() noexcept -> std::suspend_always { return {}; }
    auto yield_value(Ref value) noexcept -> std::suspend_always {
      current = std::addressof(value);
      return {};
    }
    void return_void() noexcept {}
    void unhandled_exception() { error = std::current_exception(); }
  };
  struct sentinel {};
  class iterator {
    std::coroutine_handle<promise_type> _h;

  public:
    using value_type = std::remove_cvref_t<Ref>;
    using difference_type = std::ptrdiff_t;
    explicit iterator(std::coroutine_handle<promise_type> h) : _h(h) {}
    auto operator*() const -> Ref { return static_cast<Ref>(*_h.promise().current); }
    auto operator++() -> iterator& {
      _h.resume();
      if (_h.done() && _h.promise().error) std::rethrow_exception(_h.promise().error);
      return *this;
    }
    void operator++(int) { ++*this; }
    friend bool operator==(iterator const& it, sentinel) { return it._h.done(); }
  };

  generator(generator&& o) noexcept : _h(std::exchange(o._h, {})) {}
  auto operator=(generator&& o) noexcept -> generator& {
    if (this != &o) { if (_h) _h.destroy(); _h = std::exchange(o._h, {}); }
    return *this;
  }
  generator(generator const&) = delete;
  ~generator() { if (_h) _h.destroy(); }

  auto begin() -> iterator {
    _h.resume();
    if (_h.done() && _h.promise().error) std::rethrow_exception(_h.promise().error);
    return iterator {_h};
  }
  auto end() const noexcept -> sentinel { return {}; }

private:
  explicit generator(std::coroutine_handle<promise_type> h) : _h(h) {}
  std::coroutine_handle<promise_type> _h;
};
}  // namespace pbr
