/* Minimal stand-in so the reference's make_hash.hpp compiles without Boost
 * (test infrastructure, see oracle/Makefile). */
#pragma once
#define BOOST_BINARY(x) 0b##x
