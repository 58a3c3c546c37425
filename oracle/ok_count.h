// ORACLE (test infrastructure only).  Flop-counting scalar: instantiate the oracle pipeline with
// ok::Counted instead of double to obtain the ALGORITHMIC fp64 operation count of the reference
// algorithm (with stage skipping) - SURVEY.md 8(d): add/sub/mul/div/sqrt/transcendental = 1 each.
#pragma once
#include <cmath>
#include <cstdint>

namespace ok {

struct FlopCounter {
  static uint64_t& n() {
    static thread_local uint64_t c = 0;
    return c;
  }
};

struct Counted {
  double v;
  Counted() : v(0) {}
  Counted(double x) : v(x) {}
  Counted(int x) : v(x) {}
  explicit operator double() const { return v; }
};

inline Counted operator+(Counted a, Counted b) { FlopCounter::n()++; return Counted(a.v + b.v); }
inline Counted operator-(Counted a, Counted b) { FlopCounter::n()++; return Counted(a.v - b.v); }
inline Counted operator*(Counted a, Counted b) { FlopCounter::n()++; return Counted(a.v * b.v); }
inline Counted operator/(Counted a, Counted b) { FlopCounter::n()++; return Counted(a.v / b.v); }
inline Counted operator-(Counted a) { return Counted(-a.v); }
inline Counted& operator+=(Counted& a, Counted b) { FlopCounter::n()++; a.v += b.v; return a; }
inline Counted& operator-=(Counted& a, Counted b) { FlopCounter::n()++; a.v -= b.v; return a; }
inline Counted& operator*=(Counted& a, Counted b) { FlopCounter::n()++; a.v *= b.v; return a; }
inline Counted& operator/=(Counted& a, Counted b) { FlopCounter::n()++; a.v /= b.v; return a; }
inline bool operator<(Counted a, Counted b) { return a.v < b.v; }
inline bool operator>(Counted a, Counted b) { return a.v > b.v; }
inline bool operator<=(Counted a, Counted b) { return a.v <= b.v; }
inline bool operator>=(Counted a, Counted b) { return a.v >= b.v; }
inline bool operator==(Counted a, Counted b) { return a.v == b.v; }
inline bool operator!=(Counted a, Counted b) { return a.v != b.v; }
inline Counted sqrt(Counted a) { FlopCounter::n()++; return Counted(std::sqrt(a.v)); }
inline Counted sin(Counted a) { FlopCounter::n()++; return Counted(std::sin(a.v)); }
inline Counted cos(Counted a) { FlopCounter::n()++; return Counted(std::cos(a.v)); }
inline Counted atan2(Counted a, Counted b) { FlopCounter::n()++; return Counted(std::atan2(a.v, b.v)); }
inline Counted pow(Counted a, Counted b) { FlopCounter::n()++; return Counted(std::pow(a.v, b.v)); }
inline Counted fabs(Counted a) { return Counted(std::fabs(a.v)); }
inline bool isfinite(Counted a) { return std::isfinite(a.v); }

inline double to_double(double x) { return x; }
inline double to_double(Counted x) { return x.v; }

}  // namespace ok
