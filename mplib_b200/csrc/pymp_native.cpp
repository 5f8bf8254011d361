// pymp_native.cpp -- pybind11 host module over the C-ABI of libmpb.so (include/mpb.h).
//
// The reference binds its C++ classes with pybind11 (python/pybind.cpp:16-27, python/pybind_planning_world.hpp:39-123);
// this module is the corresponding native layer of the replacement for the PER-CALL path: it owns the compiled scene
// handle and turns numpy buffers into C-ABI calls and packed result words back into numpy arrays without ctypes or
// interpreter-level bit unpacking, with the GIL released while the GPU works.  The one-off scene compiler (URDF / SRDF /
// mesh ingestion, pair enumeration) stays in mplib_b200/pymp; it hands the flat description over as numpy arrays.
//
// Build: Makefile target `native` (g++, Python.h + pybind11 headers from the venv, links libmpb.so next to it).
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/mpb.h"

namespace py = pybind11;
using f32arr = py::array_t<float, py::array::c_style | py::array::forcecast>;
using f64arr = py::array_t<double, py::array::c_style | py::array::forcecast>;
using i32arr = py::array_t<int32_t, py::array::c_style | py::array::forcecast>;

namespace {

// rc -> the exception type the reference raises at the same place: std::invalid_argument -> ValueError,
// everything else -> RuntimeError (the ASSERT macro of src/macros_utils.h:5-8)
void check(int rc) {
  if (rc == MPB_OK) return;
  const std::string msg = mpb_last_error();
  if (rc == MPB_ERR_INVALID) throw std::invalid_argument(msg);
  throw std::runtime_error("libmpb error " + std::to_string(rc) + ": " + msg);
}

py::dict stats_dict(const mpb_stats &s) {
  py::dict d;
  d["n_states"] = s.n_states; d["pairs_admitted"] = s.pairs_admitted; d["pairs_after_cull"] = s.pairs_after_cull;
  d["pairs_after_spheres"] = s.pairs_after_spheres; d["narrow_iterations"] = s.narrow_iterations; d["narrow_capped"] = s.narrow_capped;
  d["kernel_launches"] = s.kernel_launches; d["gpu_ms"] = s.gpu_ms; d["gpu_ms_cull"] = s.gpu_ms_cull; d["gpu_ms_narrow"] = s.gpu_ms_narrow;
  d["host_ms_enqueue"] = s.host_ms_enqueue; d["host_ms_total"] = s.host_ms_total; d["gpu_ms_fk"] = s.gpu_ms_fk;
  d["fk_bytes_per_state"] = s.fk_bytes_per_state; d["async_pending"] = s.async_pending;
  return d;
}

template <typename T>
const T *ptr_of(const py::dict &a, const char *name, std::vector<py::array> &keep) {
  py::array_t<T, py::array::c_style | py::array::forcecast> arr = py::cast<py::array_t<T, py::array::c_style | py::array::forcecast>>(a[name]);
  keep.push_back(arr);
  static const T dummy[4] = {};
  return arr.size() ? arr.data() : dummy;
}

class Scene {
 public:
  // `a`: the arrays of mplib_b200.scene.SceneDesc by name (the flat mpb_scene_desc)
  Scene(const py::dict &a, int device) {
    std::vector<py::array> keep;
    mpb_scene_desc d;
    std::memset(&d, 0, sizeof(d));
    auto n_of = [&](const char *name) { return (int32_t)py::cast<py::array>(a[name]).shape(0); };
    d.n_joints = n_of("joint_parent");
    d.joint_parent = ptr_of<int32_t>(a, "joint_parent", keep);
    d.joint_placement = ptr_of<double>(a, "joint_placement", keep);
    d.joint_type = ptr_of<int32_t>(a, "joint_type", keep);
    d.joint_axis = ptr_of<double>(a, "joint_axis", keep);
    d.joint_qidx = ptr_of<int32_t>(a, "joint_qidx", keep);
    d.nq_full = n_of("q_base");
    d.q_base = ptr_of<double>(a, "q_base", keep);
    d.state_dim = n_of("state_map");
    d.state_map = ptr_of<int32_t>(a, "state_map", keep);
    d.n_links = n_of("link_joint");
    d.link_joint = ptr_of<int32_t>(a, "link_joint", keep);
    d.link_placement = ptr_of<double>(a, "link_placement", keep);
    d.n_objects = n_of("obj_type");
    d.obj_type = ptr_of<int32_t>(a, "obj_type", keep);
    d.obj_link = ptr_of<int32_t>(a, "obj_link", keep);
    d.obj_local = ptr_of<double>(a, "obj_local", keep);
    d.obj_param = ptr_of<double>(a, "obj_param", keep);
    d.obj_geom = ptr_of<int32_t>(a, "obj_geom", keep);
    d.n_hulls = n_of("hull_vert_off") - 1;
    d.hull_vert_off = ptr_of<int32_t>(a, "hull_vert_off", keep);
    d.hull_verts = ptr_of<double>(a, "hull_verts", keep);
    d.hull_face_off = ptr_of<int32_t>(a, "hull_face_off", keep);
    d.hull_faces = ptr_of<int32_t>(a, "hull_faces", keep);
    d.n_meshes = n_of("mesh_vert_off") - 1;
    d.mesh_vert_off = ptr_of<int32_t>(a, "mesh_vert_off", keep);
    d.mesh_verts = ptr_of<double>(a, "mesh_verts", keep);
    d.mesh_tri_off = ptr_of<int32_t>(a, "mesh_tri_off", keep);
    d.mesh_tris = ptr_of<int32_t>(a, "mesh_tris", keep);
    d.n_octrees = n_of("oct_res");
    d.oct_res = ptr_of<double>(a, "oct_res", keep);
    d.oct_cell_off = ptr_of<int32_t>(a, "oct_cell_off", keep);
    d.oct_cells = ptr_of<int32_t>(a, "oct_cells", keep);
    d.n_pairs = n_of("pair_a");
    d.pair_a = ptr_of<int32_t>(a, "pair_a", keep);
    d.pair_b = ptr_of<int32_t>(a, "pair_b", keep);
    check(mpb_scene_create(&d, device, &h_));
    refresh_info();
  }
  ~Scene() { close(); }
  void close() {
    if (h_) {
      mpb_scene *h = h_;
      h_ = nullptr;
      py::gil_scoped_release rel;
      mpb_scene_destroy(h);
    }
  }
  uintptr_t handle() const { return reinterpret_cast<uintptr_t>(h_); }
  int n_pairs() const { return n_pairs_; }
  int state_dim() const { return state_dim_; }
  int n_links() const { return n_links_; }

  void set_q_base(const f64arr &q) {
    need();
    check(mpb_scene_set_q_base(h_, q.data(), (int32_t)q.size()));
  }
  void set_pairs(const i32arr &a, const i32arr &b) {
    need();
    if (a.size() != b.size()) throw std::invalid_argument("pair_a / pair_b differ in length");
    check(mpb_scene_set_pairs(h_, (int32_t)a.size(), a.data(), b.data()));
    refresh_info();
  }
  void update_objects(const i32arr &idx, const py::object &local, const py::object &param) {
    need();
    f64arr loc, par;
    const double *pl = nullptr, *pp = nullptr;
    if (!local.is_none()) { loc = py::cast<f64arr>(local); if (loc.size() != idx.size() * 12) throw std::invalid_argument("obj_local must be [n][12]"); pl = loc.data(); }
    if (!param.is_none()) { par = py::cast<f64arr>(param); if (par.size() != idx.size() * 4) throw std::invalid_argument("obj_param must be [n][4]"); pp = par.data(); }
    check(mpb_scene_update_objects(h_, (int32_t)idx.size(), idx.data(), pl, pp));
  }

  // q [n, D] -> (verdict bool[n], pair_hit bool[n, P] or None, stats dict or None): ValidityCheckerTpl::isValid for a batch
  py::tuple collide(const f32arr &q, bool want_pairs, bool want_stats) {
    need();
    const int64_t n = rows(q);
    const int W = (n_pairs_ + 31) / 32;
    std::vector<uint32_t> vb((size_t)(n + 31) / 32 + 1, 0u), pb(want_pairs ? (size_t)n * std::max(W, 1) : 0, 0u);
    mpb_stats st;
    int rc;
    {
      py::gil_scoped_release rel;
      rc = mpb_collide_batch(h_, q.data(), n, vb.data(), want_pairs ? pb.data() : nullptr, &st, MPB_HOST, nullptr);
    }
    check(rc);
    py::array_t<bool> verdict(n);
    bool *v = verdict.mutable_data();
    for (int64_t i = 0; i < n; i++) v[i] = (vb[(size_t)i >> 5] >> (i & 31)) & 1u;
    py::object pairs = py::none();
    if (want_pairs) {
      py::array_t<bool> ph({(py::ssize_t)n, (py::ssize_t)n_pairs_});
      bool *p = ph.mutable_data();
      for (int64_t i = 0; i < n; i++) {
        const uint32_t *row = pb.data() + (size_t)i * std::max(W, 1);
        for (int k = 0; k < n_pairs_; k++) p[(size_t)i * n_pairs_ + k] = (row[k >> 5] >> (k & 31)) & 1u;
      }
      pairs = ph;
    }
    return py::make_tuple(verdict, pairs, want_stats ? py::object(stats_dict(st)) : py::object(py::none()));
  }

  // the sparse form: (verdict bool[n], hits int64[k, 2] = (configuration, pair), sorted, stats dict)
  py::tuple collide_hits(const f32arr &q, int64_t hit_cap) {
    need();
    const int64_t n = rows(q);
    std::vector<uint32_t> vb((size_t)(n + 31) / 32 + 1, 0u);
    int64_t cap = hit_cap > 0 ? hit_cap : std::max<int64_t>(1024, 4 * n);
    std::vector<uint32_t> hits;
    uint32_t nh = 0;
    mpb_stats st;
    for (;;) {
      hits.assign((size_t)std::max<int64_t>(cap, 1) * 2, 0u);
      int rc;
      {
        py::gil_scoped_release rel;
        rc = mpb_collide_batch_hits(h_, q.data(), n, vb.data(), hits.data(), cap, &nh, &st, MPB_HOST, nullptr);
      }
      check(rc);
      if ((int64_t)nh <= cap || hit_cap > 0) break;
      cap = nh;  // the list was truncated: repeat with the reported size
    }
    const int64_t k = std::min<int64_t>(nh, cap);
    std::vector<std::pair<uint32_t, uint32_t>> h((size_t)k);
    for (int64_t i = 0; i < k; i++) h[(size_t)i] = {hits[2 * (size_t)i], hits[2 * (size_t)i + 1]};
    std::sort(h.begin(), h.end());
    py::array_t<int64_t> out({(py::ssize_t)k, (py::ssize_t)2});
    int64_t *o = out.mutable_data();
    for (int64_t i = 0; i < k; i++) { o[2 * i] = h[(size_t)i].first; o[2 * i + 1] = h[(size_t)i].second; }
    py::array_t<bool> verdict(n);
    bool *v = verdict.mutable_data();
    for (int64_t i = 0; i < n; i++) v[i] = (vb[(size_t)i >> 5] >> (i & 31)) & 1u;
    return py::make_tuple(verdict, out, stats_dict(st));
  }

  // q [n, D] -> link poses [n, n_links, 7] (x y z qw qx qy qz): computeForwardKinematics + getLinkPose
  py::array_t<float> fk(const f32arr &q) {
    need();
    const int64_t n = rows(q);
    py::array_t<float> out({(py::ssize_t)n, (py::ssize_t)n_links_, (py::ssize_t)7});
    int rc;
    {
      py::gil_scoped_release rel;
      rc = mpb_fk_batch(h_, q.data(), n, out.mutable_data(), MPB_HOST, nullptr);
    }
    check(rc);
    return out;
  }

  // q [n, D] -> (min_distance float64[n], argmin int32[n], stats): ValidityCheckerTpl::clearance for a batch
  py::tuple distance(const f32arr &q) {
    need();
    const int64_t n = rows(q);
    std::vector<float> md((size_t)std::max<int64_t>(n, 1));
    py::array_t<int32_t> am(n);
    mpb_stats st;
    int rc;
    {
      py::gil_scoped_release rel;
      rc = mpb_distance_batch(h_, q.data(), n, md.data(), am.mutable_data(), &st, MPB_HOST, nullptr);
    }
    check(rc);
    py::array_t<double> out(n);
    double *o = out.mutable_data();
    const int32_t *a = am.data();
    for (int64_t i = 0; i < n; i++) o[i] = a[i] < 0 ? std::numeric_limits<double>::max() : (double)md[(size_t)i];  // WorldDistanceResult starts at max
    return py::make_tuple(out, am, stats_dict(st));
  }

  py::tuple poll(bool wait) {
    need();
    mpb_stats st;
    std::memset(&st, 0, sizeof(st));
    int64_t done = 0, failed = 0;
    int rc;
    {
      py::gil_scoped_release rel;
      rc = mpb_scene_poll(h_, wait ? 1 : 0, &st, &done, &failed);
    }
    check(rc);
    return py::make_tuple(stats_dict(st), done, failed);
  }

 private:
  void need() const {
    if (!h_) throw std::runtime_error("the scene was closed");
  }
  int64_t rows(const f32arr &q) const {
    if (q.ndim() != 2 || q.shape(1) != state_dim_)
      throw std::runtime_error("State dimension is not correct: expected [n," + std::to_string(state_dim_) + "]");
    return (int64_t)q.shape(0);
  }
  void refresh_info() {
    int32_t np = 0, sd = 0, nl = 0, no = 0, dev = 0;
    check(mpb_scene_info(h_, &np, &sd, &nl, &no, &dev));
    n_pairs_ = np; state_dim_ = sd; n_links_ = nl;
  }
  mpb_scene *h_ = nullptr;
  int n_pairs_ = 0, state_dim_ = 0, n_links_ = 0;
};

}  // namespace

PYBIND11_MODULE(_native, m) {
  m.doc() = "pybind11 host layer of mplib_b200 over the C-ABI of libmpb.so (per-call path; see pymp_native.cpp)";
  m.def("device_count", []() { return mpb_device_count(); });
  m.def("last_error", []() { return std::string(mpb_last_error()); });
  py::class_<Scene>(m, "Scene")
      .def(py::init<const py::dict &, int>(), py::arg("arrays"), py::arg("device") = 0)
      .def("close", &Scene::close)
      .def_property_readonly("handle", &Scene::handle)
      .def_property_readonly("n_pairs", &Scene::n_pairs)
      .def_property_readonly("state_dim", &Scene::state_dim)
      .def_property_readonly("n_links", &Scene::n_links)
      .def("set_q_base", &Scene::set_q_base)
      .def("set_pairs", &Scene::set_pairs)
      .def("update_objects", &Scene::update_objects, py::arg("idx"), py::arg("obj_local") = py::none(), py::arg("obj_param") = py::none())
      .def("collide", &Scene::collide, py::arg("q"), py::arg("want_pairs") = true, py::arg("stats") = false)
      .def("collide_hits", &Scene::collide_hits, py::arg("q"), py::arg("hit_cap") = 0)
      .def("fk", &Scene::fk)
      .def("distance", &Scene::distance)
      .def("poll", &Scene::poll, py::arg("wait") = true);
}
