import sys, os, time, json
sys.path[:0]=['duckdb-fastlanes_b200','tests','.']
import numpy as np, torch, flsgpu, bench
flsgpu.init((0,))
w,img=bench.build_table_image(100_000_000, os.cpu_count())
t=flsgpu.Table(img)
n_cols=t.n_columns
is_str=[t.column(c).phys_bits==128 for c in range(n_cols)]
widths=[4 if is_str[c] else t.column(c).phys_bits//8 for c in range(n_cols)]
out_elems=t.out_elems(t.n_rowgroups)
host=[flsgpu.host_alloc(out_elems*wd) for wd in widths]
cols=[]
for c in range(n_cols):
    sc=flsgpu.ScanColumn(col=c,out_type=0,out_host=host[c],out_capacity_bytes=out_elems*widths[c])
    if is_str[c]:
        sc.string_form=1
        sc.keys_capacity_bytes=max(16,t.string_dict_keys(c)*16); sc.heap_capacity_bytes=max(16,t.string_heap_bytes(c))
        sc.keys_host=flsgpu.host_alloc(sc.keys_capacity_bytes); sc.heap_host=flsgpu.host_alloc(sc.heap_capacity_bytes)
    cols.append(sc)
dec=sum(t.n_tuples*wd for wd in widths)
ts=[]
for s in range(4):
    t.evict(); torch.cuda.synchronize()
    t0=time.perf_counter(); t.scan_projection_to_host(cols); dt=time.perf_counter()-t0
    if s: ts.append(dt)
print(os.environ.get("FLSGPU_SLAB_MB","64"), "MB slabs:", round(dec/np.mean(ts)/1e9,2), "GB/s", [round(x*1e3,1) for x in ts])
t.close(); w.close()
