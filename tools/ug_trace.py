"""Timeline of one CTA of the fused update + Gram kernel (library built with -DB200_UG_TRACE, see update_gram.cu):
per tile, when its boxes landed (update warp 0), when X_new was complete, when the Gram sweep started / ended, when the
producer re-issued the buffer.  usage: python tools/ug_trace.py trace.txt [first_tile] [ntiles]"""
import sys, collections
fn = sys.argv[1]
t0_tile = int(sys.argv[2]) if len(sys.argv) > 2 else 20
nt = int(sys.argv[3]) if len(sys.argv) > 3 else 8
ev = collections.defaultdict(list)
hdr = ""
for ln in open(fn):
    if ln.startswith("#"):
        hdr = ln.strip(); continue
    w, t, e, tile = ln.split()
    ev[int(tile)].append((int(t), int(w), int(e)))
print(hdr)
base = min(t for t, _, _ in ev[t0_tile])
def first(tile, pred):
    v = [t for t, w, e in ev[tile] if pred(w, e)]
    return (min(v) - base) / 1000.0 if v else float("nan")
def last(tile, pred):
    v = [t for t, w, e in ev[tile] if pred(w, e)]
    return (max(v) - base) / 1000.0 if v else float("nan")
print("tile | producer: group0 free .. last group free | update w0: first box .. last box | X_new done (first..last warp) | Gram start (first) .. end (last)   [us]")
for tile in range(t0_tile, t0_tile + nt):
    if tile not in ev: break
    print(f"{tile:4d} | {first(tile, lambda w, e: w == 12 and e == 5):7.2f} .. {last(tile, lambda w, e: w == 12 and 5 <= e < 10):7.2f} | "
          f"{first(tile, lambda w, e: w == 0 and e >= 10):7.2f} .. {last(tile, lambda w, e: w == 0 and e >= 10):7.2f} | "
          f"{first(tile, lambda w, e: w < 4 and e == 2):7.2f} .. {last(tile, lambda w, e: w < 4 and e == 2):7.2f} | "
          f"{first(tile, lambda w, e: 4 <= w < 12 and e == 3):7.2f} .. {last(tile, lambda w, e: 4 <= w < 12 and e == 4):7.2f}")
