// Definitions of the opaque handle types of the C ABI (include/b200comm.h),
// shared by the translation units that implement it.
#pragma once
#include "b200comm/comm.hpp"
#include "b200comm/datatype.hpp"

struct b2c_type { b200::Datatype t; };
struct b2c_comm { b200::comm::Comm c; };
struct b2c_request { b200::comm::Request r; };
