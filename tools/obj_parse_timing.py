"""Cold (fresh process) time of the host OBJ front end: python tools/obj_parse_timing.py file.obj <threads> <chunkBytes>."""
import time, os, ctypes as C, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cudapathtracer_b200 as nb
L=nb.binding.lib()
p=sys.argv[1].encode(); th=int(sys.argv[2]); ch=int(sys.argv[3])
e=(C.c_float*3)(0,0,0)
h=C.c_void_p(); t=time.time(); L.novum_obj_parse(p,C.byref(e),5,0.0,th,ch,C.byref(h)); print("threads",th,"chunk",ch,"cold total %.3f"%(time.time()-t), flush=True)
