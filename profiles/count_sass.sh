#!/bin/bash
# usage: tools_count_sass.sh <mangled-substring>   -> number of SASS instructions of matching kernels in libsnb.so
cuobjdump -sass "$(dirname "$0")/../synerfgine_b200/libsnb.so" 2>/dev/null | awk -v pat="$1" '/Function :/ {if (name!="") print n, name; name=$3; n=0; keep=index($3,pat)>0} /^ +\/\*[0-9a-f]+\*\/ / {if(keep) n++} END{print n,name}' | grep "$1"
