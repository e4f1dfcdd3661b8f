#!/bin/bash
# usage: scripts/sass_stats.sh <object> <demangled-substring>   -- registers + SASS opcode histogram of matching kernels
obj=$1; pat=$2
cuobjdump -res-usage "$obj" 2>/dev/null | grep -A1 "Function" | paste - - | sed 's/Function //' | while read -r line; do
  name=$(echo "$line" | awk '{print $1}' | sed 's/:$//')
  dem=$(echo "$name" | c++filt)
  if [[ "$dem" == *"$pat"* ]]; then
    echo "== $dem"
    echo "   $(echo "$line" | grep -oE 'REG:[0-9]+ STACK:[0-9]+')"
    cuobjdump -sass -fun "$name" "$obj" > /tmp/_k.sass
    echo "   total SASS: $(grep -cE '^\s+/\*[0-9a-f]{4}\*/' /tmp/_k.sass)"
    grep -oE "^\s+/\*[0-9a-f]{4}\*/\s+(@!?U?P[0-9T] )?[A-Z0-9_.]+" /tmp/_k.sass | awk '{print $NF}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -${3:-18} | awk '{printf "%s:%s ", $2, $1} END {print ""}'
  fi
done
