/* Minimal stand-in: boost::shared_ptr -> std::shared_ptr. */
#pragma once
#include <memory>
namespace boost {
using std::shared_ptr;
}
