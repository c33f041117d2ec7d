import os,sys,time
sys.path.insert(0,'/root/repo')
import raytracing_course_hw_b200 as rt
sc=rt.load_scene('/root/repo/tests/scenes/hw5/mixed.txt',width=1920,height=1080,samples=256)
r=rt.Renderer(); r.upload(sc)
for i in range(3):
    r.clear(); r.render(); st=r.stats(); print('hw5 1080p x256: %.1f ms %.1f Mpaths/s'%(st['kernel_ms'], st['paths']/st['kernel_ms']/1e3))
