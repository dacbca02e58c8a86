// Helpers for functions that cross the C ABI (include/b200comm.h): symbol
// visibility and exception -> return-code translation.  Every exported
// function returns 0 on success or a b200::ErrorCode; the text of the last
// failure is available per thread from b2c_last_error().
#pragma once
#include <exception>
#include <new>

#include "b200comm/error.hpp"

#define B2C_EXPORT __attribute__((visibility("default")))

#define B2C_TRY try {
#define B2C_CATCH                                        \
        return 0;                                        \
    } catch (const ::b200::Error& e) {                   \
        ::b200::set_last_error(e.what());                \
        return e.code() ? e.code() : ::b200::ERR_INTERN; \
    } catch (const std::bad_alloc&) {                    \
        ::b200::set_last_error("out of host memory");    \
        return ::b200::ERR_NOMEM;                        \
    } catch (const std::exception& e) {                  \
        ::b200::set_last_error(e.what());                \
        return ::b200::ERR_INTERN;                       \
    } catch (...) {                                      \
        ::b200::set_last_error("unknown exception");     \
        return ::b200::ERR_INTERN;                       \
    }
