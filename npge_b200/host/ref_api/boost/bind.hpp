// ref_api/boost/bind.hpp -- boost::bind / _1 for the one use on this path
// (add_opt_check(boost::bind(pattern_length, this, _1)), FragmentFinder.cpp:32), on
// std::bind, for builds without Boost.
#pragma once
#include <functional>
namespace boost {
using std::bind;
}
using std::placeholders::_1;
