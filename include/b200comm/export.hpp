// Symbol visibility of the public C++ API (the library is built with
// -fvisibility=hidden; only what is marked B200_API is exported).
#pragma once
#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif
