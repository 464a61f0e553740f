/* Minimal stand-in for BOOST_FOREACH: a C++11 range-for. */
#pragma once
#define BOOST_FOREACH(decl, range) for (decl : range)
