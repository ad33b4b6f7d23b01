// Build shim (test infrastructure): lets the reference's headers compile without Boost.
#pragma once
#define BOOST_STATIC_ASSERT(x) static_assert(x, #x)
