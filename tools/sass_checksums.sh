#!/bin/bash
# Per-kernel md5 of the SASS text (addresses and encodings stripped) of every object in a directory: two builds whose
# lists agree run the same device code.  Used to show that adding opt-in kernels left the GPU-validated kernels untouched.
# usage: sass_sums.sh objdir  -> "function md5" lines for every kernel in every object
for o in $1/*.o; do
  cuobjdump -sass "$o" 2>/dev/null | awk '
    /Function :/ { if (name != "") { close(cmd) } name=$3; cmd="md5sum | cut -c1-32 | sed \"s|^|" name " |\""; next }
    /^[ \t]+\/\*[0-9a-f]{4}\*\// { if (name != "") { line=$0; sub(/^[ \t]+\/\*[0-9a-f]+\*\/[ \t]+/, "", line); sub(/[ \t]*\/\*.*/, "", line); print line | cmd } }
    END { if (name != "") close(cmd) }'
done | sort
