#!/usr/bin/env python3
"""dev helper: print a line range of a reference file with DEBUG blocks stripped."""
import sys,re
f,a,b=sys.argv[1],int(sys.argv[2]),int(sys.argv[3])
lines=open(f,errors='replace').read().split('\n')[a-1:b]
out=[];depth=0;skip=0
i=0
n=len(lines)
while i<n:
    l=lines[i]
    s=l.strip()
    if skip:
        if s.startswith('#if'): skip+=1
        elif s.startswith('#endif'): skip-=1
        i+=1; continue
    if re.match(r'#\s*ifdef\s+DEBUG|#\s*if\s+defined\(DEBUG',s) or s.startswith('#ifdef PARALLEL_STOPWATCH'):
        skip=1; i+=1; continue
    # strip if (DEBUG_COND...) { ... } blocks
    if re.match(r'if \((DEBUG_COND|gDebugFlag)',s):
        if '{' in s:
            d=s.count('{')-s.count('}')
            i+=1
            while d>0 and i<n:
                d+=lines[i].count('{')-lines[i].count('}'); i+=1
            continue
        else:
            # single statement, skip until ';'
            while ';' not in lines[i]: i+=1
            i+=1; continue
    if s=='' : i+=1; continue
    out.append(f"{a+i}: {l}")
    i+=1
print('\n'.join(out))
