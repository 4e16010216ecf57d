#!/bin/bash
# Blackwell evidence from the built library, committed as profiles/<tag>_sass_excerpt.txt:
#   bash tools/sass_excerpt.sh > profiles/r2_sass_excerpt.txt 2>/dev/null
# (libtinyr.so itself is git-ignored; rebuild it with `python -c "import __graft_entry__ as g; g.build()"`.)
# Per kernel: the instructions that show the sm_100a mechanisms in use — UTMASTG (TMA tensor store of the RGB tile),
# FENCE.VIEW.ASYNC (fence.proxy.async before it), ACQBULK / PREEXIT (griddepcontrol.wait / launch_dependents: programmatic
# dependent launch), ATOMS.INC (the single-instruction block hand-out), STG.E.EF (streaming stores of the host delivery),
# LD/ST ...SYS (system-scope loads / stores of the peer-memory reduction) — with their addresses, then a histogram of the
# kernel's memory / special instructions.
cd "$(dirname "$0")/.."
lib=tiny_renderer_b200/libtinyr.so
echo "# cuobjdump -sass -fun <symbol> $lib   ($(cuobjdump --version | tail -1))"
echo "# $(cuobjdump -lelf $lib | tr '\n' ' ')"
echo
all=$(cuobjdump -sass $lib 2>/dev/null | grep "Function :" | awk '{print $3}')
for fn in 'k_rasterINS_15PipelineDarbouxELi128E' 'k_rasterINS_13PipelinePhongELi256E' 'k_rasterINS_14PassLightDepthELi128E' 'k_raster_viewsINS_13PipelinePhongELi128E' \
          'k_vertex_setupINS_13PipelinePhongE' 'k_frame_to_host' 'k_ao_accumulate' 'k_ao_reduce_peersILi8E'; do
  sym=$(echo "$all" | grep "$fn" | head -1)
  [ -z "$sym" ] && continue
  echo "## $(echo $sym | c++filt | sed 's/(.*//')"
  body=$(cuobjdump -sass -fun "$sym" $lib 2>/dev/null | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed 's/ *\/\* 0x[0-9a-f]* \*\/ *$//')
  echo "$body" | grep -E "UTMASTG|UTMACMDFLUSH|UTMALDG|ACQBULK|PREEXIT|ATOMS|FENCE|STG.E.EF|SYS" | awk '{ $1=$1; print "   " $0 }' | head -14
  echo "   -- $(echo "$body" | wc -l) SASS instructions; of these:" $(echo "$body" | awk '{print $2}' | sed 's/^@!*U*P[0-9T]* *//' | grep -E "^(LDG|STG|LDS|STS|ATOMS|ATOMG|REDG|RED|ATOM|UTMASTG|ACQBULK|PREEXIT|MUFU|TEX|TLD|BAR|SHFL|VOTE|LDC|LDCU|FENCE|SYNCS|MEMBAR|LD|ST)" | sed 's/\..*//' | sort | uniq -c | sort -rn | awk '{printf "%s x%s  ", $2, $1}')
  echo
done
