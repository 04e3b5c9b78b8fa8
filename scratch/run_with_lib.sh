#!/bin/bash
# usage: run_with_lib.sh <lib.so> <script args...> : runs a scratch script against an alternative build of the library
LIB=$1; shift
cp grappa_b200/libgrappa_b200.so /tmp/lib_backup.so
cp $LIB grappa_b200/libgrappa_b200.so
python "$@"; rc=$?
cp /tmp/lib_backup.so grappa_b200/libgrappa_b200.so
exit $rc
