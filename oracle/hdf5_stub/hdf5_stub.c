/* TEST INFRASTRUCTURE ONLY.  Definitions for the declarations of oracle/hdf5_stub/hdf5.h (and the two scrappie functions
 * impl/eventAligner.c calls from its fast5 path): they exist so that oracle/_ref/libsaref_ea.so loads, and abort if reached. */
#include <stdio.h>
#include <stdlib.h>
/* the HDF5 and scrappie symbols eventAligner.c refers to: defined so that the library loads; never reached */
#define SA_STUB(name) int name() { fprintf(stderr, "oracle/hdf5_stub: " #name " called (HDF5 / scrappie are not built)\n"); abort(); return 0; }
SA_STUB(H5Tinsert) SA_STUB(H5Tclose) SA_STUB(H5Dread) SA_STUB(H5Tset_size) SA_STUB(H5Gclose) SA_STUB(H5Dclose) SA_STUB(H5Sclose)
SA_STUB(H5Eset_auto) SA_STUB(H5Aread) SA_STUB(H5Aclose) SA_STUB(H5Tset_cset) SA_STUB(H5Fclose) SA_STUB(H5Dwrite) SA_STUB(H5Dvlen_reclaim)
SA_STUB(H5Gget_objinfo) SA_STUB(H5Eset_auto2) SA_STUB(H5Eset_auto1) SA_STUB(H5Gopen) SA_STUB(H5Tcopy) SA_STUB(H5Dopen) SA_STUB(H5Dget_space)
SA_STUB(H5Aopen) SA_STUB(H5Tcreate) SA_STUB(H5Screate_simple) SA_STUB(H5Gcreate2) SA_STUB(H5Fopen) SA_STUB(H5Dopen2) SA_STUB(H5Dget_type)
SA_STUB(H5Dcreate2) SA_STUB(H5Aget_type) SA_STUB(H5Gopen2) SA_STUB(H5Aopen_name) SA_STUB(H5Gcreate) SA_STUB(H5Dcreate)
SA_STUB(H5Sget_simple_extent_dims) SA_STUB(H5Lget_name_by_idx) SA_STUB(H5Lexists) SA_STUB(H5Tis_variable_str) SA_STUB(H5Aexists)
SA_STUB(H5Tget_size) SA_STUB(H5Tget_cset) SA_STUB(H5Aget_storage_size) SA_STUB(detect_events) SA_STUB(trim_and_segment_raw)
long long H5T_NATIVE_DOUBLE, H5T_NATIVE_FLOAT, H5T_NATIVE_UINT64, H5T_NATIVE_INT, H5T_NATIVE_UINT, H5T_NATIVE_INT64, H5T_NATIVE_UINT32,
    H5T_NATIVE_LONG, H5T_NATIVE_ULONG, H5T_NATIVE_INT16, H5T_C_S1, H5T_STD_I16LE, H5T_NATIVE_SHORT, H5T_NATIVE_CHAR;

