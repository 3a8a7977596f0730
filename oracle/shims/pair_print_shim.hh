// Test infrastructure only.  g++ 13 two-phase lookup rejects SBWT/include/sbwt/stdlib_printing.hh
// because the vector printer uses operator<<(pair) before it is declared; forward-declare it.
#pragma once
#include <ostream>
#include <utility>
namespace sbwt { template <class S, class T> std::ostream& operator<<(std::ostream&, const std::pair<S, T>&); }
