import importlib, sys
sys.path.insert(0,'/root/repo'); import bench
pkg = importlib.import_module("svt-av1-1_b200")
frames = bench.synth_sequence(9, seed=1)
ctx = pkg.MeContext(bench.W, bench.H, 9)
for i in range(9): ctx.upload(i, frames[i])
def run(tag, over, tls=None):
    specs=[]
    for poc, tl, r0, r1 in [j for j in bench.gop_jobs() if j[0] <= 8]:
        if tls is not None and tl not in tls: continue
        p = pkg.default_params(bench.W, bench.H, 2, tl, 3, 1 if tl < 3 else 0)
        p.same_ref_poc = 1 if r0 == r1 else 0
        for k,v in over.items(): setattr(p,k,v)
        specs.append((poc, r0, r1, p))
    jobs, keep, _ = ctx.make_jobs(specs)      # `keep`: the jobs point into this host array
    run.keep = keep
    ctx.set_profiling(True)
    for _ in range(2): ctx.motion_estimate_async(jobs)
    ctx.kernel_times(reset=True)
    for _ in range(5): ctx.motion_estimate_async(jobs)
    kt = ctx.kernel_times(reset=True)
    ctx.synchronize()
    print(tag, len(specs), 'jobs: hme ms/job', round(kt['hme'][0]/5/len(specs),4), 'fullpel', round(kt['fullpel'][0]/5/len(specs),4), 'subpel', round(kt['subpel'][0]/5/len(specs),4))
run('all', {})
run('no hme (center only)', {'enable_hme_flag':0})
run('L0 only', {'enable_hme_level1':0,'enable_hme_level2':0})
run('L0+L1', {'enable_hme_level2':0})
run('L1+L2', {'enable_hme_level0':0})
run('L2 only', {'enable_hme_level0':0,'enable_hme_level1':0})
for tl in range(4): run('all tl%d'%tl, {}, [tl])
run('nonref off zero check', {'is_used_as_reference':0})
