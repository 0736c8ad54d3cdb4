// oracle/compat/ccutils/macros.hpp — TEST INFRASTRUCTURE (oracle). ccutils is an external,
// un-vendored dependency of the reference; dp.cpp:30 includes this header but uses
// nothing from it by name.
#pragma once
