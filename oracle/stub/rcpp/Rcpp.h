/* Rcpp.h -- TEST INFRASTRUCTURE: a stand-in for the few pieces of Rcpp that the reference's GFLLibrary*Impl.cpp /
 * GFLLibrary.hpp use (Rcpp::List::create with _["name"] = value, containsElementNamed, operator[], Rcpp::as<>,
 * Rcpp::stop, Rcpp::Rcout), so that oracle/Makefile can compile those files where they lie under /root/reference
 * (R and Rcpp are absent from this image). Nothing of Rcpp is copied: this is a minimal re-declaration of its API. */
#pragma once
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

namespace Rcpp {

struct Any {                       /* what a list element can hold on this path */
    double d = 0.0;
    std::vector<double> vd;
    bool is_vec = false;
    Any() {}
    Any(double x) : d(x) {}
    Any(int x) : d(x) {}
    Any(unsigned x) : d(x) {}
    Any(const std::vector<double> &v) : vd(v), is_vec(true) {}
};
struct NamedValue { std::string name; Any val; };
struct NamedPlaceholder {
    struct Proxy {
        std::string n;
        template <class T> NamedValue operator=(const T &v) const { return NamedValue{n, Any(v)}; }
    };
    Proxy operator[](const char *n) const { return Proxy{n}; }
};
static const NamedPlaceholder _ = NamedPlaceholder();

class List {
public:
    template <class... A> static List create(const A &...a) { List l; (l.items.push_back(a), ...); return l; }
    bool containsElementNamed(const char *n) const { for (auto &it : items) if (it.name == n) return true; return false; }
    const Any &operator[](const char *n) const {
        for (auto &it : items) if (it.name == n) return it.val;
        throw std::runtime_error(std::string("no list element named ") + n);
    }
    std::vector<NamedValue> items;
};

template <class T> struct AsImpl { static T get(const Any &a) { return static_cast<T>(a.d); } };
template <> struct AsImpl<std::vector<double> > { static std::vector<double> get(const Any &a) { return a.vd; } };
template <class T> T as(const Any &a) { return AsImpl<T>::get(a); }

[[noreturn]] inline void stop(const std::string &msg) { throw std::runtime_error(msg); }
static std::ostream &Rcout = std::cout;

}  // namespace Rcpp
