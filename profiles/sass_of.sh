#!/bin/bash
# usage: sass_of.sh lib.so kernel_name_substring  -> one instruction per line (address, text)
cuobjdump -sass "$1" | awk -v pat="$2" '/Function :/ {on = index($0, pat) > 0} on && /^ +\/\*[0-9a-f]+\*\/ +[A-Z@!]/ {sub(/^ +\/\*/, ""); sub(/\*\/ +/, " "); sub(/ *;.*/, ""); print}'
