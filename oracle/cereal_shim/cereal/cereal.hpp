// Minimal stand-in for USCiLab/cereal 1.3.2 (NOT vendored in the reference; it is
// fetched by CPM at configure time, lib/hibf/cmake/package-lock.cmake:11-17).
//
// TEST INFRASTRUCTURE ONLY: this header exists so that the reference's own
// sources under /root/reference compile offline into oracle/_ref/.  It is our
// own code, written against cereal's documented *binary archive* format:
//   * arithmetic types: raw native little-endian bytes, no padding
//   * std::string / std::vector<T>: uint64 size, then elements
//     (arithmetic T as one contiguous blob)
//   * std::array<T,N>: elements only, no size
//   * classes: member save/load preferred, then member serialize, then
//     non-member save/load found by ADL (hibf/cereal/path.hpp)
//   * no class-version tags (nothing in the reference registers one)
// Everything lives in this one header; the other cereal/*.hpp files of the
// shim only include it.
#pragma once

#include <array>
#include <cstdint>
#include <cstring>
#include <istream>
#include <ostream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#define CEREAL_SERIALIZE_FUNCTION_NAME serialize
#define CEREAL_LOAD_FUNCTION_NAME load
#define CEREAL_SAVE_FUNCTION_NAME save
#define CEREAL_NVP(x) ::cereal::make_nvp(#x, x)

namespace cereal
{

struct Exception : public std::runtime_error
{
    explicit Exception(std::string const & what) : std::runtime_error(what) {}
    explicit Exception(char const * what) : std::runtime_error(what) {}
};

namespace detail
{
struct OutputArchiveBase
{};
struct InputArchiveBase
{};
} // namespace detail

namespace traits
{
struct TextArchive
{};
} // namespace traits

enum class specialization
{
    member_serialize,
    member_load_save,
    member_load_save_minimal,
    non_member_serialize,
    non_member_load_save,
    non_member_load_save_minimal
};

// The reference specialises this for bit_vector; the shim's dispatch already
// prefers member load/save, so the primary template only has to exist.
template <class Archive, class T, specialization S>
struct specialize : public std::false_type
{};

template <class T>
struct BinaryData
{
    T data;
    uint64_t size;
};

template <class T>
inline BinaryData<T> binary_data(T && data, size_t size)
{
    return {std::forward<T>(data), static_cast<uint64_t>(size)};
}

template <class T>
struct NameValuePair
{
    T & value;
};

template <class T>
inline NameValuePair<T> make_nvp(char const *, T & value)
{
    return {value};
}
template <class T>
inline NameValuePair<T> make_nvp(std::string const &, T & value)
{
    return {value};
}

template <class Base>
struct base_class
{
    template <class Derived>
    base_class(Derived const * derived) : base_ptr(const_cast<Base *>(static_cast<Base const *>(derived)))
    {}
    Base * base_ptr;
};

class access
{
public:
    template <class A, class T>
    static auto member_serialize(A & a, T & t) -> decltype(t.serialize(a))
    {
        return t.serialize(a);
    }
    template <class A, class T>
    static auto member_save(A & a, T const & t) -> decltype(t.save(a))
    {
        return t.save(a);
    }
    template <class A, class T>
    static auto member_load(A & a, T & t) -> decltype(t.load(a))
    {
        return t.load(a);
    }
};

namespace shim
{
template <class A, class T>
concept has_member_save = requires(A & a, T const & t) { access::member_save(a, t); };
template <class A, class T>
concept has_member_load = requires(A & a, T & t) { access::member_load(a, t); };
template <class A, class T>
concept has_member_serialize = requires(A & a, T & t) { access::member_serialize(a, t); };
template <class T>
concept raw_pod = std::is_arithmetic_v<T> || std::is_enum_v<T>;

template <class T>
struct is_std_vector : std::false_type
{};
template <class T, class Al>
struct is_std_vector<std::vector<T, Al>> : std::true_type
{};
template <class T>
struct is_std_array : std::false_type
{};
template <class T, size_t N>
struct is_std_array<std::array<T, N>> : std::true_type
{};
template <class T>
struct is_binary_data : std::false_type
{};
template <class T>
struct is_binary_data<BinaryData<T>> : std::true_type
{};
template <class T>
struct is_nvp : std::false_type
{};
template <class T>
struct is_nvp<NameValuePair<T>> : std::true_type
{};
template <class T>
struct is_base_class : std::false_type
{};
template <class T>
struct is_base_class<base_class<T>> : std::true_type
{};
} // namespace shim

class BinaryOutputArchive : public detail::OutputArchiveBase
{
public:
    explicit BinaryOutputArchive(std::ostream & stream) : os(stream) {}

    template <class... Ts>
    BinaryOutputArchive & operator()(Ts &&... args)
    {
        (process(std::forward<Ts>(args)), ...);
        return *this;
    }

    void write_raw(void const * data, uint64_t bytes)
    {
        os.write(static_cast<char const *>(data), static_cast<std::streamsize>(bytes));
        if (!os)
            throw Exception("Failed to write " + std::to_string(bytes) + " bytes to output stream!");
    }

private:
    std::ostream & os;

    template <class T>
    void process(T const & value)
    {
        using U = std::remove_cvref_t<T>;
        if constexpr (shim::is_nvp<U>::value)
            process(value.value);
        else if constexpr (shim::is_binary_data<U>::value)
            write_raw(value.data, value.size);
        else if constexpr (shim::is_base_class<U>::value)
            process(*value.base_ptr);
        else if constexpr (shim::raw_pod<U>)
            write_raw(&value, sizeof(U));
        else if constexpr (shim::has_member_save<BinaryOutputArchive, U>)
            access::member_save(*this, value);
        else if constexpr (shim::has_member_serialize<BinaryOutputArchive, U>)
            access::member_serialize(*this, const_cast<U &>(value));
        else if constexpr (std::is_same_v<U, std::string>)
        {
            uint64_t n = value.size();
            write_raw(&n, 8);
            write_raw(value.data(), n);
        }
        else if constexpr (shim::is_std_vector<U>::value)
        {
            using E = typename U::value_type;
            uint64_t n = value.size();
            write_raw(&n, 8);
            if constexpr (std::is_same_v<E, bool>)
                for (bool b : value)
                    process(b);
            else if constexpr (shim::raw_pod<E>)
                write_raw(value.data(), n * sizeof(E));
            else
                for (auto const & e : value)
                    process(e);
        }
        else if constexpr (shim::is_std_array<U>::value)
        {
            using E = typename U::value_type;
            if constexpr (shim::raw_pod<E>)
                write_raw(value.data(), value.size() * sizeof(E));
            else
                for (auto const & e : value)
                    process(e);
        }
        else
            CEREAL_SAVE_FUNCTION_NAME(*this, value); // non-member, ADL
    }
};

class BinaryInputArchive : public detail::InputArchiveBase
{
public:
    explicit BinaryInputArchive(std::istream & stream) : is(stream) {}

    template <class... Ts>
    BinaryInputArchive & operator()(Ts &&... args)
    {
        (process(std::forward<Ts>(args)), ...);
        return *this;
    }

    void read_raw(void * data, uint64_t bytes)
    {
        is.read(static_cast<char *>(data), static_cast<std::streamsize>(bytes));
        if (static_cast<uint64_t>(is.gcount()) != bytes)
            throw Exception("Failed to read " + std::to_string(bytes) + " bytes from input stream! Read "
                            + std::to_string(is.gcount()));
    }

private:
    std::istream & is;

    template <class T>
    void process(T && value)
    {
        using U = std::remove_cvref_t<T>;
        if constexpr (shim::is_nvp<U>::value)
            process(value.value);
        else if constexpr (shim::is_binary_data<U>::value)
            read_raw(const_cast<void *>(static_cast<void const *>(value.data)), value.size);
        else if constexpr (shim::is_base_class<U>::value)
            process(*value.base_ptr);
        else if constexpr (shim::raw_pod<U>)
            read_raw(&value, sizeof(U));
        else if constexpr (shim::has_member_load<BinaryInputArchive, U>)
            access::member_load(*this, value);
        else if constexpr (shim::has_member_serialize<BinaryInputArchive, U>)
            access::member_serialize(*this, value);
        else if constexpr (std::is_same_v<U, std::string>)
        {
            uint64_t n{};
            read_raw(&n, 8);
            value.resize(n);
            read_raw(value.data(), n);
        }
        else if constexpr (shim::is_std_vector<U>::value)
        {
            using E = typename U::value_type;
            uint64_t n{};
            read_raw(&n, 8);
            value.resize(n);
            if constexpr (std::is_same_v<E, bool>)
                for (uint64_t i = 0; i < n; ++i)
                {
                    bool b{};
                    process(b);
                    value[i] = b;
                }
            else if constexpr (shim::raw_pod<E>)
                read_raw(value.data(), n * sizeof(E));
            else
                for (auto & e : value)
                    process(e);
        }
        else if constexpr (shim::is_std_array<U>::value)
        {
            using E = typename U::value_type;
            if constexpr (shim::raw_pod<E>)
                read_raw(value.data(), value.size() * sizeof(E));
            else
                for (auto & e : value)
                    process(e);
        }
        else
            CEREAL_LOAD_FUNCTION_NAME(*this, value); // non-member, ADL
    }
};

// hibf's config (de)serialises itself as JSON into layout files; TetRex never
// reads or writes a layout file, so these archives only have to compile.
class JSONOutputArchive : public detail::OutputArchiveBase, public traits::TextArchive
{
public:
    explicit JSONOutputArchive(std::ostream &) {}
    template <class... Ts>
    JSONOutputArchive & operator()(Ts &&...)
    {
        return *this;
    }
};

class JSONInputArchive : public detail::InputArchiveBase, public traits::TextArchive
{
public:
    explicit JSONInputArchive(std::istream &) {}
    template <class... Ts>
    JSONInputArchive & operator()(Ts &&...)
    {
        throw Exception("JSON archives are not supported by the oracle's cereal stand-in");
    }
};

} // namespace cereal
