// <boost/outcome/result.hpp> stand-in: value-or-error with the members the reference uses (common/adt/result.hpp).
// Test infrastructure (oracle/_ref build); Boost is not installed on this image.
#ifndef SHIM_BOOST_OUTCOME_RESULT
#define SHIM_BOOST_OUTCOME_RESULT
#include <cstdlib>
#include <utility>
#include <variant>
namespace boost { namespace outcome_v2 {
namespace policy { struct terminate {}; struct all_narrow {}; }
template <typename T, typename E, typename P = policy::terminate>
class result {
    public:
        result(const T& v) : d(std::in_place_index<0>, v) {}
        result(T&& v) : d(std::in_place_index<0>, std::move(v)) {}
        result(const E& e) : d(std::in_place_index<1>, e) {}
        bool has_value() const noexcept { return d.index() == 0; }
        bool has_error() const noexcept { return d.index() == 1; }
        explicit operator bool() const noexcept { return has_value(); }
        T& value() { if (!has_value()) std::abort(); return std::get<0>(d); }
        const T& value() const { if (!has_value()) std::abort(); return std::get<0>(d); }
        E& error() { if (!has_error()) std::abort(); return std::get<1>(d); }
        const E& error() const { if (!has_error()) std::abort(); return std::get<1>(d); }
        T& assume_value() { return std::get<0>(d); }
        const T& assume_value() const { return std::get<0>(d); }
        E& assume_error() { return std::get<1>(d); }
        const E& assume_error() const { return std::get<1>(d); }
    private:
        std::variant<T, E> d;
};
template <typename T, typename E> using checked = result<T, E, policy::terminate>;
template <typename T, typename E> using unchecked = result<T, E, policy::all_narrow>;
}}  // namespace boost::outcome_v2
#endif
