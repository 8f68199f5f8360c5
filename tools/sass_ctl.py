#!/usr/bin/env python
"""Developer tool: list a kernel's SASS (cuobjdump -sass output) with the scheduling fields of each instruction decoded from
the second encoding word: stall count (s), yield, the scoreboard the instruction sets when its result arrives (W0..W5) or
when its operands have been read (R0..R5), and the scoreboards it waits for (wait:..).

    cuobjdump -sass -fun <mangled name> file.o > k.sass ; python tools/sass_ctl.py k.sass [first_addr_hex last_addr_hex]

Used to find out which loads an instruction really waits for (all global loads of the aggregation sweep share one scoreboard).
"""
import re,sys
lines=open(sys.argv[1]).read().split('\n')
out=[]
i=0
pat=re.compile(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);\s*/\* (0x[0-9a-f]{16}) \*/')
pat2=re.compile(r'/\* (0x[0-9a-f]{16}) \*/')
while i<len(lines):
    m=pat.search(lines[i])
    if m and i+1<len(lines):
        m2=pat2.search(lines[i+1])
        if m2:
            hi=int(m2.group(1),16)
            stall=(hi>>41)&0xf; y=(hi>>45)&1; wb=(hi>>46)&7; rb=(hi>>49)&7; wm=(hi>>52)&0x3f
            out.append((m.group(1), m.group(2).strip(), stall,y,wb,rb,wm))
            i+=2; continue
    i+=1
lo=int(sys.argv[2],16) if len(sys.argv)>2 else 0
hi_=int(sys.argv[3],16) if len(sys.argv)>3 else 1<<30
for a,ins,stall,y,wb,rb,wm in out:
    if lo<=int(a,16)<=hi_:
        print(a, ins[:70].ljust(70), 's%-2d'%stall, 'Y' if y else ' ', 'W%d'%wb if wb!=7 else '  ', 'R%d'%rb if rb!=7 else '  ', 'wait:'+''.join(str(b) for b in range(6) if wm>>b&1) if wm else '')
