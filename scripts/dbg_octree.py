import numpy as np, sys
sys.path.insert(0,'.')
from tests.workloads import c3_world
from tests.common import random_q
from oracle import oracle as O
w = c3_world(primitives=False, mesh=False, cloud_points=20000, attach=True)
q = random_q(3000, 22)
c = w._sync(); desc = c["dev"].desc
v, ph, st = c["dev"].collide(q, want_pairs=True, stats=True)
qd = q.astype(np.float32).astype(np.float64)
ov, oph = O.collide_batch(desc, qd, threads=8)
mism = np.argwhere(ph != oph)
poses = O.object_poses(desc, qd)
for i,p in mism:
    sd = O.pair_signed_distance(desc, qd[i], int(p))
    a,b = desc.pair_a[p], desc.pair_b[p]
    print(i,p,desc.object_names[a],desc.object_names[b],'gpu',ph[i,p],'orc',oph[i,p],'sd',sd)
    if abs(sd) >= 1e-4:
        # exact check: hull vertices in world vs voxel boxes
        h = desc.obj_geom[a]; V = desc.hull_verts[desc.hull_vert_off[h]:desc.hull_vert_off[h+1]]
        F = desc.hull_faces[desc.hull_face_off[h]:desc.hull_face_off[h+1]]
        T = poses[i,a]; Vw = V@T[:,:3].T + T[:,3]
        res = desc.oct_res[0]; cells = desc.oct_cells
        cen = (cells+0.5)*res
        lo, hi = Vw.min(0), Vw.max(0)
        near = np.all((cen+res/2 >= lo-1e-3)&(cen-res/2 <= hi+1e-3),axis=1)
        print('  near voxels', near.sum(), 'hull aabb', lo, hi)
        # SAT per near voxel
        E = np.array([Vw[F[:,(k+1)%3]]-Vw[F[:,k]] for k in range(3)]).reshape(-1,3)
        N = np.cross(Vw[F[:,1]]-Vw[F[:,0]], Vw[F[:,2]]-Vw[F[:,0]])
        axes = [N, np.eye(3)] + [np.cross(E, np.eye(3)[k][None,:]) for k in range(3)]
        axes = np.concatenate(axes); nrm = np.linalg.norm(axes,axis=1); axes = axes[nrm>1e-12]/nrm[nrm>1e-12][:,None]
        best=-1e9
        for cc in cen[near]:
            corners = cc + (np.array(np.meshgrid([-1,1],[-1,1],[-1,1])).reshape(3,-1).T)*res/2
            ph_ = Vw@axes.T; pb = corners@axes.T
            gap = np.maximum(pb.min(0)-ph_.max(0), ph_.min(0)-pb.max(0)).max()
            best=max(best,-gap)
        print('  max penetration over near voxels (SAT, >0 means intersect):', best)
