//! Safe layer over `b200rt-sys`.  One type per reference concept:
//! `SceneBuilder` = what `Hitable::flatten` / `Material::flatten` / `Texture::flatten` write into,
//! `Scene` = `World` (world.rs:15-29), `Scene::render*` = `Camera::capture_image` (camera.rs:66-128),
//! `Scene::render_sppm` = `SPPMIntegrator::new` + `capture_image` (main.rs:51-53).
//! NOT compiled in the image this repository was built in (no Rust toolchain there): see ../README.md.
use b200rt_sys as sys;
use std::ffi::{CStr, CString};
use std::os::raw::c_int;
use std::ptr;

pub use sys::{b200rt_camera as Camera, b200rt_hit as Hit, b200rt_multi_stats as MultiStats, b200rt_params as Params,
              b200rt_sppm_params as SppmParams, b200rt_sppm_stats as SppmStats, b200rt_stats as Stats};

/// Every call of the C ABI returns a status; nothing panics or aborts across it (the reference panics instead:
/// vec3.rs:88, bvh.rs:43,57, transform.rs:146, mesh.rs:158).
#[derive(Debug, Clone)]
pub struct Error { pub status: i32, pub message: String }
impl std::fmt::Display for Error {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result { write!(f, "b200rt status {}: {}", self.status, self.message) }
}
impl std::error::Error for Error {}
pub type Result<T> = std::result::Result<T, Error>;

unsafe fn text(p: *const std::os::raw::c_char) -> String {
    if p.is_null() { String::new() } else { CStr::from_ptr(p).to_string_lossy().into_owned() }
}

/// One device, one stream.  Not `Sync`: one context per host thread, like the C ABI says.
pub struct Context { raw: *mut sys::b200rt_ctx, pub device: i32 }
unsafe impl Send for Context {}
impl Context {
    /// Fails with `B200RT_ERR_CUDA` when there is no device: the library has no CPU fallback.
    pub fn new(device: i32) -> Result<Self> {
        let mut raw = ptr::null_mut();
        let rc = unsafe { sys::b200rt_ctx_create(device as c_int, &mut raw) };
        if rc != 0 { return Err(Error { status: rc, message: unsafe { text(sys::b200rt_global_error()) } }); }
        Ok(Self { raw, device })
    }
    pub fn device_count() -> i32 {
        let mut n: c_int = 0;
        unsafe { sys::b200rt_device_count(&mut n) };
        n as i32
    }
    fn check(&self, rc: c_int) -> Result<()> {
        if rc == 0 { Ok(()) } else { Err(Error { status: rc, message: unsafe { text(sys::b200rt_last_error(self.raw)) } }) }
    }
    /// `From<Vec3> for Rgb<u8>` (vec3.rs:223-231) over a float4 accumulation image.
    pub fn tonemap(&self, accum: &[f32], width: i32, height: i32) -> Result<Vec<u8>> {
        assert_eq!(accum.len(), 4 * width as usize * height as usize);
        let mut rgb = vec![0u8; 3 * width as usize * height as usize];
        self.check(unsafe { sys::b200rt_tonemap_rgb8(self.raw, accum.as_ptr(), width, height, rgb.as_mut_ptr()) })?;
        Ok(rgb)
    }
}
impl Drop for Context { fn drop(&mut self) { unsafe { sys::b200rt_ctx_destroy(self.raw); } } }

/// `RgbImage::save` (main.rs:54).
pub fn write_png(path: &str, rgb8: &[u8], width: i32, height: i32) -> Result<()> {
    let c = CString::new(path).map_err(|_| Error { status: sys::B200RT_ERR_INVALID, message: "path contains NUL".into() })?;
    let rc = unsafe { sys::b200rt_write_png(c.as_ptr(), rgb8.as_ptr(), width, height) };
    if rc == 0 { Ok(()) } else { Err(Error { status: rc, message: format!("cannot write {path}") }) }
}

/// `Camera::new` (camera.rs:24-55).
pub fn camera(look_from: [f64; 3], look_at: [f64; 3], vup: [f64; 3], vfov_deg: f64, aspect: f64, aperture: f64, focus_dist: f64) -> Result<Camera> {
    let mut cam = Camera::default();
    let rc = unsafe { sys::b200rt_camera_make(look_from.as_ptr(), look_at.as_ptr(), vup.as_ptr(), vfov_deg, aspect, aperture, focus_dist, 0.0, 0.0, &mut cam) };
    if rc == 0 { Ok(cam) } else { Err(Error { status: rc, message: "degenerate camera".into() }) }
}

/// The flat tables a scene is lowered into.  The reference's `Hitable` / `Material` / `Texture` objects gain one
/// `flatten(&self, &mut SceneBuilder) -> i32` method each that calls these.
pub struct SceneBuilder { raw: *mut sys::b200rt_builder }
impl SceneBuilder {
    pub fn new() -> Result<Self> {
        let mut raw = ptr::null_mut();
        let rc = unsafe { sys::b200rt_builder_create(&mut raw) };
        if rc != 0 { return Err(Error { status: rc, message: "builder_create".into() }); }
        Ok(Self { raw })
    }
    /// add_* return the new id (>= 0) or a negative status
    fn id(&self, rc: c_int) -> Result<i32> {
        if rc >= 0 { Ok(rc as i32) } else { Err(Error { status: -rc, message: unsafe { text(sys::b200rt_builder_error(self.raw)) } }) }
    }
    fn ok(&self, rc: c_int) -> Result<()> {
        if rc == 0 { Ok(()) } else { Err(Error { status: rc, message: unsafe { text(sys::b200rt_builder_error(self.raw)) } }) }
    }
    // Texture (material.rs:48-84)
    pub fn texture_constant(&mut self, c: [f64; 3]) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_texture_constant(self.raw, c[0], c[1], c[2]) }) }
    pub fn texture_checker(&mut self, c0: [f64; 3], c1: [f64; 3]) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_texture_checker(self.raw, c0.as_ptr(), c1.as_ptr()) }) }
    pub fn texture_image(&mut self, rgb8: &[u8], width: i32, height: i32) -> Result<i32> {
        assert_eq!(rgb8.len(), 3 * width as usize * height as usize);
        self.id(unsafe { sys::b200rt_builder_texture_image(self.raw, rgb8.as_ptr(), width, height) })
    }
    pub fn texture_noise(&mut self, scale: f64) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_texture_noise(self.raw, scale) }) }
    // Material (material.rs:86-231): param = Metal fuzz / Dielectric ir
    pub fn lambertian(&mut self, texture: i32) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_material(self.raw, sys::B200RT_MAT_LAMBERTIAN, texture, 0.0) }) }
    pub fn metal(&mut self, texture: i32, fuzz: f64) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_material(self.raw, sys::B200RT_MAT_METAL, texture, fuzz) }) }
    pub fn dielectric(&mut self, texture: i32, ir: f64) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_material(self.raw, sys::B200RT_MAT_DIELECTRIC, texture, ir) }) }
    pub fn diffuse_light(&mut self, texture: i32) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_material(self.raw, sys::B200RT_MAT_DIFFUSE_LIGHT, texture, 0.0) }) }
    pub fn isotropic(&mut self, texture: i32) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_material(self.raw, sys::B200RT_MAT_ISOTROPIC, texture, 0.0) }) }
    // Hitable
    /// Sphere::new (sphere.rs:10-14)
    pub fn sphere(&mut self, center: [f64; 3], radius: f64, material: i32) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_sphere(self.raw, center.as_ptr(), radius, material) }) }
    pub fn moving_sphere(&mut self, c0: [f64; 3], c1: [f64; 3], t0: f64, t1: f64, radius: f64, material: i32) -> Result<i32> {
        self.id(unsafe { sys::b200rt_builder_moving_sphere(self.raw, c0.as_ptr(), c1.as_ptr(), t0, t1, radius, material) })
    }
    /// XY / XZ / YZRectangle (rectangle.rs): axis 2 / 1 / 0, k = the fixed coordinate
    pub fn rect(&mut self, axis: i32, a0: f64, b0: f64, a1: f64, b1: f64, k: f64, material: i32) -> Result<i32> {
        self.id(unsafe { sys::b200rt_builder_rect(self.raw, axis, a0, b0, a1, b1, k, material) })
    }
    /// Cube::new (cube.rs:16-61)
    pub fn cube(&mut self, min: [f64; 3], max: [f64; 3], material: i32) -> Result<i32> { self.id(unsafe { sys::b200rt_builder_cube(self.raw, min.as_ptr(), max.as_ptr(), material) }) }
    /// Triangle::new (mesh.rs:19-54) with its three vertex normals
    pub fn triangle(&mut self, p: [[f64; 3]; 3], n: [[f64; 3]; 3], material: i32) -> Result<i32> {
        self.id(unsafe { sys::b200rt_builder_triangle(self.raw, p[0].as_ptr(), p[1].as_ptr(), p[2].as_ptr(), n[0].as_ptr(), n[1].as_ptr(), n[2].as_ptr(), material) })
    }
    /// Mesh::load_obj (mesh.rs:148-199)
    pub fn mesh_obj(&mut self, path: &str, material: i32, make_normals: bool, subdivide: i32) -> Result<i32> {
        let c = CString::new(path).map_err(|_| Error { status: sys::B200RT_ERR_INVALID, message: "path contains NUL".into() })?;
        self.id(unsafe { sys::b200rt_builder_mesh_obj(self.raw, c.as_ptr(), material, make_normals as i32, subdivide) })
    }
    /// ConstantMedium::new (medium.rs:15-23) over a Sphere / Cube boundary
    pub fn medium_sphere(&mut self, center: [f64; 3], radius: f64, density: f64, phase: i32) -> Result<i32> {
        self.id(unsafe { sys::b200rt_builder_medium_sphere(self.raw, center.as_ptr(), radius, density, phase) })
    }
    pub fn medium_cube(&mut self, min: [f64; 3], max: [f64; 3], density: f64, phase: i32) -> Result<i32> {
        self.id(unsafe { sys::b200rt_builder_medium_cube(self.raw, min.as_ptr(), max.as_ptr(), density, phase) })
    }
    /// Transform::new(rotate_in_degree, scale, translate, obj) (transform.rs:17-148): everything added inside `f` carries it
    pub fn with_transform<T>(&mut self, rotate_deg: [f64; 3], scale: [f64; 3], translate: [f64; 3], f: impl FnOnce(&mut Self) -> Result<T>) -> Result<T> {
        self.ok(unsafe { sys::b200rt_builder_push_transform(self.raw, rotate_deg.as_ptr(), scale.as_ptr(), translate.as_ptr()) })?;
        let r = f(self);
        self.ok(unsafe { sys::b200rt_builder_pop_transform(self.raw) })?;
        r
    }
    /// the `lights` argument of World::new (world.rs:15-25): primitives the light PDF / the photon emitter sample
    pub fn mark_light(&mut self, prim: i32) -> Result<()> { self.ok(unsafe { sys::b200rt_builder_mark_light(self.raw, prim) }) }
    /// with the light's own `scale` (light.rs:71,131), relative to `SppmParams::light_scale`
    pub fn mark_light_scaled(&mut self, prim: i32, scale: f64) -> Result<()> { self.ok(unsafe { sys::b200rt_builder_mark_light_scaled(self.raw, prim, scale) }) }
    /// `Vec<Arc<dyn Hitable>>::hit` instead of the BVH (hit.rs:56-67)
    pub fn flat_list(&mut self) -> Result<()> { self.ok(unsafe { sys::b200rt_builder_set_flags(self.raw, sys::B200RT_SCENE_FLAT_LIST) }) }
    /// data/scene_*.json / *.yaml (Appendix B of SURVEY.md)
    pub fn load_json(&mut self, text: &str) -> Result<Camera> {
        let mut cam = Camera::default();
        self.ok(unsafe { sys::b200rt_builder_load_json(self.raw, text.as_ptr().cast(), text.len() as u64, &mut cam) })?;
        Ok(cam)
    }
    pub fn load_yaml(&mut self, text: &str) -> Result<Camera> {
        let mut cam = Camera::default();
        self.ok(unsafe { sys::b200rt_builder_load_yaml(self.raw, text.as_ptr().cast(), text.len() as u64, &mut cam) })?;
        Ok(cam)
    }
    /// Host only: the surface-area estimate of the BVH `Scene::new` would build (node steps, primitive tests per random line) and
    /// the scene's dense box, if it has one (`b200rt_scene_bvh_quality`; no counterpart in the reference).
    pub fn bvh_quality(&mut self) -> Result<(f32, f32, Option<([f32; 3], [f32; 3])>)> {
        let mut desc: sys::b200rt_scene_desc = unsafe { std::mem::zeroed() };
        self.ok(unsafe { sys::b200rt_builder_finish(self.raw, &mut desc) })?;
        let (mut nodes, mut prims, mut has) = (0f32, 0f32, 0i32);
        let (mut lo, mut hi) = ([0f32; 3], [0f32; 3]);
        self.ok(unsafe { sys::b200rt_scene_bvh_quality(&desc, &mut nodes, &mut prims, &mut has, lo.as_mut_ptr(), hi.as_mut_ptr()) })?;
        Ok((nodes, prims, if has != 0 { Some((lo, hi)) } else { None }))
    }
}
impl Drop for SceneBuilder { fn drop(&mut self) { unsafe { sys::b200rt_builder_destroy(self.raw); } } }

/// `World` (world.rs:15-29) on one device: immutable once created, like `Arc<World>`.
pub struct Scene<'c> { raw: *mut sys::b200rt_scene, ctx: &'c Context }
impl<'c> Scene<'c> {
    /// `World::new`: validates the tables, bakes the transforms, builds the wide BVH (replaces `BVHNode::new`,
    /// bvh.rs:60-83) and uploads everything in one copy.
    pub fn new(ctx: &'c Context, builder: &mut SceneBuilder) -> Result<Self> {
        let mut desc: sys::b200rt_scene_desc = unsafe { std::mem::zeroed() };
        builder.ok(unsafe { sys::b200rt_builder_finish(builder.raw, &mut desc) })?;
        let mut raw = ptr::null_mut();
        ctx.check(unsafe { sys::b200rt_scene_create(ctx.raw, &desc, &mut raw) })?;
        Ok(Self { raw, ctx })
    }
    /// `World::hit` (world.rs:27-29) over a batch: rows are (o.xyz, time) and (d.xyz, t_max).
    pub fn hit_batch(&self, o_time: &[[f32; 4]], d_tmax: &[[f32; 4]], t_min: f32) -> Result<Vec<Hit>> {
        assert_eq!(o_time.len(), d_tmax.len());
        let mut out = vec![Hit::default(); o_time.len()];
        self.ctx.check(unsafe {
            sys::b200rt_intersect_batch(self.raw, o_time.as_ptr().cast(), d_tmax.as_ptr().cast(), ptr::null(), o_time.len() as u64, t_min, out.as_mut_ptr())
        })?;
        Ok(out)
    }
    /// `Camera::capture_image` (camera.rs:66-128): float4 sums + sample count per pixel.
    pub fn render(&self, cam: &Camera, params: &Params) -> Result<(Vec<f32>, Stats)> {
        let mut accum = vec![0f32; 4 * params.width as usize * params.height as usize];
        let mut st = Stats::default();
        self.ctx.check(unsafe { sys::b200rt_render(self.raw, cam, params, accum.as_mut_ptr(), &mut st) })?;
        Ok((accum, st))
    }
    /// `SPPMIntegrator::new(world)` + `capture_image(integrator)` (main.rs:51-53).
    pub fn render_sppm(&self, cam: &Camera, params: &Params, sppm: &SppmParams) -> Result<(Vec<f32>, SppmStats)> {
        let mut accum = vec![0f32; 4 * params.width as usize * params.height as usize];
        let mut st = SppmStats::default();
        self.ctx.check(unsafe { sys::b200rt_render_sppm(self.raw, cam, params, sppm, accum.as_mut_ptr(), ptr::null_mut(), &mut st) })?;
        Ok((accum, st))
    }
}
impl Drop for Scene<'_> { fn drop(&mut self) { unsafe { sys::b200rt_scene_destroy(self.raw); } } }

/// `capture_image` sharded by sample index over the devices of `scenes` (the same scene created on one `Context` per
/// GPU), one peer-memory or NCCL sum-reduce onto the first device.
pub fn render_multi(scenes: &[&Scene<'_>], cam: &Camera, params: &Params) -> Result<(Vec<f32>, MultiStats)> {
    assert!(!scenes.is_empty());
    let raws: Vec<*mut sys::b200rt_scene> = scenes.iter().map(|s| s.raw).collect();
    let mut accum = vec![0f32; 4 * params.width as usize * params.height as usize];
    let mut ms = MultiStats::default();
    scenes[0].ctx.check(unsafe { sys::b200rt_render_multi(raws.len() as i32, raws.as_ptr(), cam, params, accum.as_mut_ptr(), ptr::null_mut(), &mut ms) })?;
    Ok((accum, ms))
}

/// The SPPM integrator (`Scene::render_sppm`) over the devices of `scenes`: photons sharded by id, the photon maps gathered inside
/// each device's grid build over NVLink peer mappings, pixel statistics sharded by pixel, the final pass by sample index
/// (`b200rt_render_sppm_multi`).  Needs peer access between the devices.
pub fn render_sppm_multi(scenes: &[&Scene<'_>], cam: &Camera, params: &Params, sppm: &SppmParams) -> Result<(Vec<f32>, SppmStats)> {
    assert!(!scenes.is_empty());
    let raws: Vec<*mut sys::b200rt_scene> = scenes.iter().map(|s| s.raw).collect();
    let mut accum = vec![0f32; 4 * params.width as usize * params.height as usize];
    let mut st = SppmStats::default();
    scenes[0].ctx.check(unsafe { sys::b200rt_render_sppm_multi(raws.len() as i32, raws.as_ptr(), cam, params, sppm, accum.as_mut_ptr(), ptr::null_mut(), &mut st) })?;
    Ok((accum, st))
}

/// The reference's defaults: 800x800 (main.rs:33-46), 256 spp (camera.rs:73), depth 50 and t_min 0.001 (photon_mapper.rs:333-339).
pub fn default_params() -> Params {
    Params { width: 800, height: 800, spp_begin: 0, spp_count: 256, max_depth: 50, integrator: sys::B200RT_INTEGRATOR_PATH,
             pipeline: sys::B200RT_PIPELINE_WAVEFRONT, background_mode: sys::B200RT_BACKGROUND_CONSTANT, t_min: 0.001, background: [0.0; 3],
             flags: 0, pad: 0, seed: 1 }
}
/// photon_mapper.rs:17-19, 148-149; scene.rs:26-32 (light scale)
pub fn default_sppm() -> SppmParams {
    SppmParams { iterations: 50, photons_per_iter: 500_000, k_global: 100, k_caustic: 50, alpha: 0.7, light_scale: 1e6 }
}
