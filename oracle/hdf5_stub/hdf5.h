/* TEST INFRASTRUCTURE ONLY.  Declarations-only stand-in for <hdf5.h>, so that the reference's impl/eventAligner.c compiles
 * UNMODIFIED where HDF5 is not installed (this image).  Only its fast5-free function adaptive_banded_simple_event_align2
 * is ever called (oracle/ea_harness.c); the HDF5 entry points are declared without prototypes and never resolved. */
#ifndef SA_HDF5_STUB_H
#define SA_HDF5_STUB_H
#include <stddef.h>
#include <stdint.h>
typedef int64_t hid_t;
typedef int herr_t;
typedef int htri_t;
typedef unsigned long long hsize_t;
typedef long long hssize_t;
typedef int H5T_cset_t;
typedef int H5T_class_t;
typedef struct { unsigned long fileno[2]; unsigned long objno[2]; unsigned nlink; int type; long mtime; size_t linklen; } H5G_stat_t;
typedef struct { void *p; size_t len; } hvl_t;
#define H5P_DEFAULT 0
#define H5S_ALL 0
#define H5F_ACC_RDWR 1u
#define H5F_ACC_RDONLY 0u
#define H5T_VARIABLE ((size_t) -1)
#define H5T_COMPOUND 6
#define H5T_CSET_ASCII 0
#define H5T_CSET_UTF8 1
#define H5_ITER_INC 0
#define H5_INDEX_NAME 0
#define H5G_GROUP 0
#define H5G_DATASET 1
#define H5E_BEGIN_TRY {
#define H5E_END_TRY }
#define HOFFSET(S, M) (offsetof(S, M))
extern hid_t H5T_NATIVE_DOUBLE, H5T_NATIVE_FLOAT, H5T_NATIVE_UINT64, H5T_NATIVE_INT, H5T_NATIVE_UINT, H5T_NATIVE_INT64,
    H5T_NATIVE_UINT32, H5T_NATIVE_LONG, H5T_NATIVE_ULONG, H5T_NATIVE_INT16, H5T_C_S1, H5T_STD_I16LE, H5T_NATIVE_SHORT, H5T_NATIVE_CHAR;
herr_t H5Tinsert(), H5Tclose(), H5Dread(), H5Tset_size(), H5Gclose(), H5Dclose(), H5Sclose(), H5Eset_auto(), H5Aread(), H5Aclose(),
    H5Tset_cset(), H5Fclose(), H5Dwrite(), H5Dvlen_reclaim(), H5Gget_objinfo(), H5Eset_auto2(), H5Eset_auto1();
hid_t H5Gopen(), H5Tcopy(), H5Dopen(), H5Dget_space(), H5Aopen(), H5Tcreate(), H5Screate_simple(), H5Gcreate2(), H5Fopen(),
    H5Dopen2(), H5Dget_type(), H5Dcreate2(), H5Aget_type(), H5Gopen2(), H5Aopen_name(), H5Gcreate(), H5Dcreate();
int H5Sget_simple_extent_dims();
long long H5Lget_name_by_idx();
htri_t H5Lexists(), H5Tis_variable_str(), H5Aexists();
size_t H5Tget_size();
H5T_cset_t H5Tget_cset();
hsize_t H5Aget_storage_size();
#endif
