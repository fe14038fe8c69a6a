/* Stub: the reference includes <hdf5.h> unconditionally in one lightcone
 * header; no HDF5 is installed and none of it is reached by the hydro path. */
#ifndef REF_STUB_HDF5_H
#define REF_STUB_HDF5_H
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
typedef long hid_t;
typedef int herr_t;
typedef unsigned long long hsize_t;
typedef int htri_t;
#endif
