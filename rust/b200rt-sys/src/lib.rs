//! Raw bindings to `libb200rt.so` — one declaration per entry point of `include/b200rt.h` (ABI version 1).
//! Written by hand so the crate builds without bindgen; with `--features bindgen` the same names come from the header.
//! NOT compiled in the image this repository was built in (no Rust toolchain there): see ../README.md.
#![allow(non_camel_case_types, non_upper_case_globals)]

#[cfg(feature = "bindgen")]
include!(concat!(env!("OUT_DIR"), "/bindings.rs"));

#[cfg(not(feature = "bindgen"))]
mod manual {
    use core::ffi::{c_char, c_int, c_void};

    pub const B200RT_ABI_VERSION: c_int = 1;
    // status codes
    pub const B200RT_OK: c_int = 0;
    pub const B200RT_ERR_INVALID: c_int = 1;
    pub const B200RT_ERR_CUDA: c_int = 2; // CUDA runtime error, or no device: there is no CPU fallback
    pub const B200RT_ERR_NOMEM: c_int = 3;
    pub const B200RT_ERR_UNSUPPORTED: c_int = 4;
    pub const B200RT_ERR_IO: c_int = 5;
    pub const B200RT_ERR_PARSE: c_int = 6;
    // enum b200rt_material_type (material.rs:89-231)
    pub const B200RT_MAT_LAMBERTIAN: i32 = 0;
    pub const B200RT_MAT_METAL: i32 = 1;
    pub const B200RT_MAT_DIELECTRIC: i32 = 2;
    pub const B200RT_MAT_DIFFUSE_LIGHT: i32 = 3;
    pub const B200RT_MAT_ISOTROPIC: i32 = 4;
    // enum b200rt_integrator / b200rt_pipeline
    pub const B200RT_INTEGRATOR_PATH: i32 = 0;
    pub const B200RT_INTEGRATOR_MIXTURE: i32 = 1;
    pub const B200RT_INTEGRATOR_NEE: i32 = 2;
    pub const B200RT_PIPELINE_WAVEFRONT: i32 = 0;
    pub const B200RT_PIPELINE_MEGAKERNEL: i32 = 1;
    pub const B200RT_BACKGROUND_CONSTANT: i32 = 0;
    pub const B200RT_BACKGROUND_SKY: i32 = 1;
    pub const B200RT_FLAG_LIGHT_SCATTERS: u32 = 1;
    pub const B200RT_SCENE_FLAT_LIST: u32 = 1;
    pub const B200RT_REDUCE_NONE: i32 = 0;
    pub const B200RT_REDUCE_NCCL: i32 = 1;
    pub const B200RT_REDUCE_PEER: i32 = 2;

    #[repr(C)] pub struct b200rt_ctx { _private: [u8; 0] }
    #[repr(C)] pub struct b200rt_scene { _private: [u8; 0] }
    #[repr(C)] pub struct b200rt_builder { _private: [u8; 0] }

    /// The flat scene tables (`World::new`'s arguments, world.rs:15-25).  Filled by `b200rt_builder_finish`; the table
    /// element types are only ever handled through the builder from Rust, so the pointers stay untyped here.
    #[repr(C)]
    #[derive(Clone, Copy)]
    pub struct b200rt_scene_desc {
        pub n_prims: u32, pub n_tri_normals: u32, pub n_materials: u32, pub n_textures: u32,
        pub n_images: u32, pub n_instances: u32, pub n_lights: u32, pub flags: u32,
        pub image_bytes: u64,
        pub prims: *const c_void, pub tri_normals: *const f32, pub materials: *const c_void, pub textures: *const c_void,
        pub images: *const c_void, pub image_rgb8: *const u8, pub instances: *const c_void, pub light_prims: *const i32,
        pub perlin: *const c_void, pub light_scales: *const f32,
    }

    /// camera.rs:11-21 plus a shutter interval.
    #[repr(C)]
    #[derive(Clone, Copy, Default, Debug)]
    pub struct b200rt_camera {
        pub origin: [f32; 3], pub lower_left_corner: [f32; 3], pub horizontal: [f32; 3], pub vertical: [f32; 3],
        pub u: [f32; 3], pub v: [f32; 3], pub w: [f32; 3], pub lens_radius: f32, pub time0: f32, pub time1: f32,
    }

    /// The compile-time constants of main.rs:34-46, camera.rs:73, photon_mapper.rs:334-335.
    #[repr(C)]
    #[derive(Clone, Copy, Default, Debug)]
    pub struct b200rt_params {
        pub width: i32, pub height: i32, pub spp_begin: i32, pub spp_count: i32, pub max_depth: i32,
        pub integrator: i32, pub pipeline: i32, pub background_mode: i32, pub t_min: f32, pub background: [f32; 3],
        pub flags: u32, pub pad: u32, pub seed: u64,
    }

    #[repr(C)]
    #[derive(Clone, Copy, Default, Debug)]
    pub struct b200rt_stats {
        pub paths: u64, pub extend_rays: u64, pub shadow_rays: u64, pub kernel_launches: u64,
        pub ms_total: f32, pub ms_generate: f32, pub ms_extend: f32, pub ms_shade: f32, pub ms_shadow: f32, pub ms_other: f32,
        pub wavefront_iterations: u32, pub pipelines: u32,
        pub ms_generate_sum: f32, pub ms_extend_sum: f32, pub ms_shade_sum: f32, pub ms_shadow_sum: f32,
    }

    /// `HitRecord` (hit.rs:7-39) of one ray; prim = index into the scene's primitive list or -1.
    #[repr(C)]
    #[derive(Clone, Copy, Default, Debug)]
    pub struct b200rt_hit { pub t: f32, pub prim: i32, pub nx: f32, pub ny: f32, pub nz: f32, pub front_face: i32, pub u: f32, pub v: f32 }

    #[repr(C)]
    #[derive(Clone, Copy, Default, Debug)]
    pub struct b200rt_multi_stats {
        pub paths: u64, pub extend_rays: u64, pub shadow_rays: u64,
        pub ms_render_max: f32, pub ms_reduce: f32, pub ms_wall: f32, pub reduce_path: i32, pub n_dev: i32, pub pad: u32,
    }

    /// photon_mapper.rs:17, 51-52, 148-149
    #[repr(C)]
    #[derive(Clone, Copy, Debug)]
    pub struct b200rt_sppm_params { pub iterations: i32, pub photons_per_iter: i32, pub k_global: i32, pub k_caustic: i32, pub alpha: f64, pub light_scale: f64 }

    #[repr(C)]
    #[derive(Clone, Copy, Default, Debug)]
    pub struct b200rt_sppm_stats {
        pub global_photons: u64, pub caustic_photons: u64, pub photon_rays: u64, pub eye_rays: u64,
        pub ms_total: f32, pub ms_photons: f32, pub ms_grid: f32, pub ms_eye: f32, pub ms_final: f32, pub pad: u32,
    }

    extern "C" {
        // ---- context
        pub fn b200rt_abi_version() -> c_int;
        pub fn b200rt_device_count(count: *mut c_int) -> c_int;
        pub fn b200rt_ctx_create(device: c_int, out: *mut *mut b200rt_ctx) -> c_int;
        pub fn b200rt_ctx_destroy(ctx: *mut b200rt_ctx) -> c_int;
        pub fn b200rt_ctx_set_stream(ctx: *mut b200rt_ctx, cuda_stream: *mut c_void) -> c_int;
        pub fn b200rt_last_error(ctx: *const b200rt_ctx) -> *const c_char;
        pub fn b200rt_global_error() -> *const c_char;
        pub fn b200rt_ctx_device_info(ctx: *const b200rt_ctx, sm_count: *mut i32, clock_khz: *mut i32, l2_bytes: *mut i64, name: *mut c_char, name_len: c_int) -> c_int;
        // ---- World::new (world.rs:15-25)
        pub fn b200rt_scene_create(ctx: *mut b200rt_ctx, desc: *const b200rt_scene_desc, out: *mut *mut b200rt_scene) -> c_int;
        pub fn b200rt_scene_destroy(scene: *mut b200rt_scene) -> c_int;
        pub fn b200rt_scene_info(scene: *const b200rt_scene, h2d_bytes: *mut u64, bvh_nodes: *mut u32, device_prims: *mut u32) -> c_int;
    pub fn b200rt_scene_bvh_quality(desc: *const b200rt_scene_desc, node_steps: *mut f32, prim_tests: *mut f32, has_dense_box: *mut i32, dense_lo: *mut f32, dense_hi: *mut f32) -> c_int;
    pub fn b200rt_scene_compile_stats(desc: *const b200rt_scene_desc, bvh_nodes: *mut u32, bvh4_depth: *mut u32, stack_needed: *mut u32, stack_capacity: *mut u32) -> c_int;
    pub fn b200rt_debug_eval(scene: *mut b200rt_scene, cam: *const b200rt_camera, params: *const b200rt_params, kind: i32, n: u64, in_host: *const f32, out_host: *mut f32) -> c_int;
        // ---- World::hit over a batch (world.rs:27-29)
        pub fn b200rt_intersect_batch(scene: *mut b200rt_scene, o_time: *const f32, d_tmax: *const f32, medium_seed: *const u32, n: u64, t_min: f32, out: *mut b200rt_hit) -> c_int;
        pub fn b200rt_intersect_batch_device(scene: *mut b200rt_scene, o_time: *const f32, d_tmax: *const f32, medium_seed: *const u32, n: u64, t_min: f32, out: *mut b200rt_hit) -> c_int;
        // ---- Camera::capture_image (camera.rs:66-128)
        pub fn b200rt_render_device(scene: *mut b200rt_scene, cam: *const b200rt_camera, params: *const b200rt_params, accum_dev: *mut f32, stats: *mut b200rt_stats) -> c_int;
    pub fn b200rt_render_device_static(scene: *mut b200rt_scene, cam: *const b200rt_camera, params: *const b200rt_params, accum_dev: *mut f32, stats: *mut b200rt_stats) -> c_int;
        pub fn b200rt_render(scene: *mut b200rt_scene, cam: *const b200rt_camera, params: *const b200rt_params, accum_host: *mut f32, stats: *mut b200rt_stats) -> c_int;
        pub fn b200rt_render_multi(n_dev: i32, scenes: *const *mut b200rt_scene, cam: *const b200rt_camera, params: *const b200rt_params, accum_host: *mut f32,
                                   stats_per_device: *mut b200rt_stats, multi: *mut b200rt_multi_stats) -> c_int;
        // ---- SPPMIntegrator (integrator/photon_mapper.rs)
        pub fn b200rt_render_sppm(scene: *mut b200rt_scene, cam: *const b200rt_camera, params: *const b200rt_params, sppm: *const b200rt_sppm_params,
                                  accum_host: *mut f32, pixel_stats_host: *mut f32, stats: *mut b200rt_sppm_stats) -> c_int;
    pub fn b200rt_render_sppm_multi(n_dev: i32, scenes: *const *mut b200rt_scene, cam: *const b200rt_camera, params: *const b200rt_params, sppm: *const b200rt_sppm_params,
                                  accum_host: *mut f32, pixel_stats_host: *mut f32, stats: *mut b200rt_sppm_stats) -> c_int;
        // ---- From<Vec3> for Rgb<u8> (vec3.rs:223-231), RgbImage::save (main.rs:54), Camera::new (camera.rs:24-55)
        pub fn b200rt_tonemap_rgb8(ctx: *mut b200rt_ctx, accum_host: *const f32, width: i32, height: i32, rgb_out: *mut u8) -> c_int;
        pub fn b200rt_tonemap_rgb8_device(ctx: *mut b200rt_ctx, accum_dev: *const f32, width: i32, height: i32, rgb_dev: *mut u8) -> c_int;
        pub fn b200rt_write_png(path: *const c_char, rgb8: *const u8, width: i32, height: i32) -> c_int;
        pub fn b200rt_camera_make(look_from: *const f64, look_at: *const f64, vup: *const f64, vfov_deg: f64, aspect: f64, aperture: f64, focus_dist: f64,
                                  time0: f64, time1: f64, out: *mut b200rt_camera) -> c_int;
        pub fn b200rt_measure_fp32_peak(ctx: *mut b200rt_ctx, tflops: *mut f64, sm_mhz_effective: *mut f64) -> c_int;
        // ---- scene builder: the `flatten` target (host only).  add_* return the new id (>= 0) or a negative status.
        pub fn b200rt_builder_create(out: *mut *mut b200rt_builder) -> c_int;
        pub fn b200rt_builder_destroy(b: *mut b200rt_builder) -> c_int;
        pub fn b200rt_builder_error(b: *const b200rt_builder) -> *const c_char;
        pub fn b200rt_builder_texture_constant(b: *mut b200rt_builder, r: f64, g: f64, bl: f64) -> c_int;
        pub fn b200rt_builder_texture_checker(b: *mut b200rt_builder, c0: *const f64, c1: *const f64) -> c_int;
        pub fn b200rt_builder_texture_image(b: *mut b200rt_builder, rgb8: *const u8, width: i32, height: i32) -> c_int;
        pub fn b200rt_builder_texture_noise(b: *mut b200rt_builder, scale: f64) -> c_int;
        pub fn b200rt_builder_material(b: *mut b200rt_builder, type_: i32, texture: i32, param: f64) -> c_int;
        pub fn b200rt_builder_push_transform(b: *mut b200rt_builder, rotate_deg: *const f64, scale: *const f64, translate: *const f64) -> c_int;
        pub fn b200rt_builder_pop_transform(b: *mut b200rt_builder) -> c_int;
        pub fn b200rt_builder_sphere(b: *mut b200rt_builder, center: *const f64, radius: f64, material: i32) -> c_int;
        pub fn b200rt_builder_moving_sphere(b: *mut b200rt_builder, center0: *const f64, center1: *const f64, time0: f64, time1: f64, radius: f64, material: i32) -> c_int;
        pub fn b200rt_builder_rect(b: *mut b200rt_builder, axis: i32, a0: f64, b0: f64, a1: f64, b1: f64, k: f64, material: i32) -> c_int;
        pub fn b200rt_builder_cube(b: *mut b200rt_builder, box_min: *const f64, box_max: *const f64, material: i32) -> c_int;
        pub fn b200rt_builder_triangle(b: *mut b200rt_builder, pa: *const f64, pb: *const f64, pc: *const f64, na: *const f64, nb: *const f64, nc: *const f64, material: i32) -> c_int;
        pub fn b200rt_builder_mesh_obj(b: *mut b200rt_builder, path: *const c_char, material: i32, make_normals: i32, subdivide: i32) -> c_int;
        pub fn b200rt_builder_medium_sphere(b: *mut b200rt_builder, center: *const f64, radius: f64, density: f64, phase_material: i32) -> c_int;
        pub fn b200rt_builder_medium_cube(b: *mut b200rt_builder, box_min: *const f64, box_max: *const f64, density: f64, phase_material: i32) -> c_int;
        pub fn b200rt_builder_mark_light(b: *mut b200rt_builder, prim: i32) -> c_int;
        pub fn b200rt_builder_mark_light_scaled(b: *mut b200rt_builder, prim: i32, scale: f64) -> c_int;
        pub fn b200rt_builder_set_flags(b: *mut b200rt_builder, flags: u32) -> c_int;
        pub fn b200rt_builder_seed_perlin(b: *mut b200rt_builder, seed: u64) -> c_int;
        pub fn b200rt_builder_load_json(b: *mut b200rt_builder, text: *const c_char, len: u64, cam_out: *mut b200rt_camera) -> c_int;
        pub fn b200rt_builder_load_yaml(b: *mut b200rt_builder, text: *const c_char, len: u64, cam_out: *mut b200rt_camera) -> c_int;
        pub fn b200rt_builder_make_scene(b: *mut b200rt_builder, name: *const c_char, seed: u64, aux_path: *const c_char, earth_rgb8: *const u8, earth_w: i32, earth_h: i32,
                                         aspect: f64, cam_out: *mut b200rt_camera, params_out: *mut b200rt_params) -> c_int;
        pub fn b200rt_builder_finish(b: *mut b200rt_builder, out: *mut b200rt_scene_desc) -> c_int;
    }
}
#[cfg(not(feature = "bindgen"))]
pub use manual::*;
