/* Minimal stand-in: boost::make_shared -> std::make_shared. */
#pragma once
#include <memory>
namespace boost {
using std::make_shared;
}
