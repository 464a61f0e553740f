// ref_api/Block.hpp -- stands where the reference's header of this name would be when the
// processors are compiled with -DNPGE_B200_REAL_HEADERS outside an NPGE checkout: the
// facade carries the reference's own signatures too (string rules, 3-argument add_gopt,
// add_opt_check), so this branch is compiled and run by the tests.
#pragma once
#include "../npge_facade.hpp"
