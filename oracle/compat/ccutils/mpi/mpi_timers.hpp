// oracle/compat/ccutils/mpi/mpi_timers.hpp — TEST INFRASTRUCTURE (oracle).
// Re-creation of the ccutils timer macros from their call sites in the reference
// (dp.cpp:69-70,102-104,258-263; fsdp.cpp:61-66,86-161,384-397): a timer is a global
// std::vector<float> __timer_vals_<name> of elapsed SECONDS, one sample per START/STOP pair.
#pragma once
#include <mpi.h>
#include <vector>

#define CCUTILS_MPI_TIMER_DEF(name)                                                                \
    std::vector<float> __timer_vals_##name;                                                        \
    double __timer_start_##name = 0.0;
#define CCUTILS_MPI_TIMER_START(name) __timer_start_##name = MPI_Wtime();
#define CCUTILS_MPI_TIMER_STOP(name) __timer_vals_##name.push_back((float)(MPI_Wtime() - __timer_start_##name));
