/* Stand-in for the site's <mpi.h>, only so that the reference's src/util/shared.h compiles in the engine check
 * (integration/Makefile): the check calls no MPI function.  A deployment uses its real MPI. */
#pragma once
typedef int MPI_Comm;
typedef int MPI_Datatype;
#define MPI_COMM_WORLD 0
#define MPI_FLOAT 4
#define MPI_DOUBLE 8
