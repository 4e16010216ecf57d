//! tiny_renderer_sys — `extern "C"` declarations for include/tinyr.h and a safe `Scene` with the method names of
//! the reference's `scene::Scene` (src/scene.rs:47-151), so `app.rs` can swap `crate::scene::Scene` for it.
//! UNBUILT ON THIS BOX (no cargo/rustc in the build image); kept in step with the header by hand.
#![allow(non_camel_case_types)]

use std::ffi::{c_char, c_int, c_void, CStr, CString};

#[repr(C)]
pub struct tr_scene { _private: [u8; 0] }

#[repr(C)]
pub struct tr_multi { _private: [u8; 0] }

#[repr(C)]
#[derive(Clone, Copy)]
pub struct tr_image { pub rgb8_topdown: *const u8, pub width: u32, pub height: u32 }

#[repr(C)]
pub struct tr_scene_desc {
    pub width: u32, pub height: u32,
    pub positions: *const f32, pub n_positions: u32,
    pub uvs: *const f32, pub n_uvs: u32,
    pub normals: *const f32, pub n_normals: u32,
    pub tri_idx: *const u32, pub n_triangles: u32,
    pub texture: tr_image, pub normal_map: tr_image, pub normal_map_tangent: tr_image, pub specular_map: tr_image,
    pub pipeline_name: *const c_char,
    pub device: i32,
    pub ring: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct tr_view { pub look_from: [f32; 3], pub look_at: [f32; 3], pub up: [f32; 3], pub light_direction: [f32; 3] }

extern "C" {
    pub fn tr_scene_create(out: *mut *mut tr_scene, desc: *const tr_scene_desc) -> c_int;
    pub fn tr_scene_create_from_dir(out: *mut *mut tr_scene, asset_dir: *const c_char, pipeline: *const c_char,
                                    width: u32, height: u32, device: i32, ring: u32) -> c_int;
    pub fn tr_scene_destroy(s: *mut tr_scene);
    pub fn tr_scene_clear(s: *mut tr_scene) -> c_int;
    pub fn tr_scene_set_light_direction(s: *mut tr_scene, dir: *const f32) -> c_int;
    pub fn tr_scene_set_camera(s: *mut tr_scene, from: *const f32, at: *const f32, up: *const f32) -> c_int;
    pub fn tr_scene_render(s: *mut tr_scene) -> c_int;
    pub fn tr_scene_get_frame_buffer(s: *mut tr_scene, dst_rgb8_topdown: *mut u8) -> c_int;
    pub fn tr_scene_get_z_buffer(s: *mut tr_scene, dst_rgb8_topdown: *mut u8) -> c_int;
    pub fn tr_scene_get_shadow_buffer(s: *mut tr_scene, dst_rgb8_topdown: *mut u8) -> c_int;
    pub fn tr_scene_read_visibility(s: *mut tr_scene, dst_tri_id: *mut u32) -> c_int;
    pub fn tr_scene_read_depth(s: *mut tr_scene, dst: *mut f32) -> c_int;
    pub fn tr_scene_render_views(s: *mut tr_scene, views: *const tr_view, n: u32, checksums: *mut u64, frames: *mut u8) -> c_int;
    pub fn tr_scene_render_ao(s: *mut tr_scene, lights_xyz: *const f32, n: u32, reset: c_int, accum_host: *mut u32) -> c_int;
    pub fn tr_scene_ao_device_buffer(s: *mut tr_scene, ptr: *mut *mut c_void, n: *mut usize) -> c_int;
    pub fn tr_scene_diagnostics(s: *mut tr_scene, out: *mut u64) -> c_int;
    pub fn tr_host_alloc(bytes: usize) -> *mut c_void;
    pub fn tr_host_free(p: *mut c_void);
    pub fn tr_host_invalidate(p: *mut c_void, bytes: usize) -> c_int;
    pub fn tr_multi_create_from_dir(out: *mut *mut tr_multi, asset_dir: *const c_char, pipeline: *const c_char, width: u32, height: u32,
                                    devices: *const i32, n_devices: u32, ring: u32) -> c_int;
    pub fn tr_multi_destroy(m: *mut tr_multi);
    pub fn tr_multi_set_camera(m: *mut tr_multi, from: *const f32, at: *const f32, up: *const f32) -> c_int;
    pub fn tr_multi_render_views(m: *mut tr_multi, views: *const tr_view, n: u32, checksums: *mut u64, frames: *mut u8) -> c_int;
    pub fn tr_multi_render_ao(m: *mut tr_multi, lights_xyz: *const f32, n: u32, accum_host: *mut u32) -> c_int;
    pub fn tr_write_tga(path: *const c_char, rgb8_topdown: *const u8, width: u32, height: u32, rle: c_int) -> c_int;
    pub fn tr_last_error() -> *const c_char;
}

/// Error carrying the tr_status code and the message of tr_last_error().  The reference panics in these places
/// (shader.rs:108, scene.rs:218, shader.rs:224,277,278,921); the backend reports instead.
#[derive(Debug)]
pub struct Error { pub status: i32, pub message: String }

fn check(rc: c_int) -> Result<(), Error> {
    if rc == 0 { return Ok(()); }
    let message = unsafe { CStr::from_ptr(tr_last_error()) }.to_string_lossy().into_owned();
    Err(Error { status: rc, message })
}

/// Same surface as the reference's `Scene` (src/scene.rs); not `Sync`, like the `&mut self` original.
pub struct Scene { h: *mut tr_scene, width: u32, height: u32 }

impl Scene {
    /// Scene::new (scene.rs:47-56) with the asset loading of app.rs:87-131 done by the backend's loaders.
    pub fn from_dir(width: u32, height: u32, asset_dir: &str, shader_pipeline_name: &str) -> Result<Scene, Error> {
        let (dir, name) = (CString::new(asset_dir).unwrap(), CString::new(shader_pipeline_name).unwrap());
        let mut h = std::ptr::null_mut();
        check(unsafe { tr_scene_create_from_dir(&mut h, dir.as_ptr(), name.as_ptr(), width, height, 0, 1) })?;
        Ok(Scene { h, width, height })
    }
    /// Scene::new with caller-owned arrays (positions xyz, uvs uv, normals xyz, 9 indices per polygon, 4 RGB8 maps).
    pub fn new(desc: &tr_scene_desc) -> Result<Scene, Error> {
        let mut h = std::ptr::null_mut();
        check(unsafe { tr_scene_create(&mut h, desc) })?;
        Ok(Scene { h, width: desc.width, height: desc.height })
    }
    pub fn clear(&mut self) -> Result<(), Error> { check(unsafe { tr_scene_clear(self.h) }) }                       // scene.rs:128
    pub fn set_light_direction(&mut self, d: [f32; 3]) -> Result<(), Error> {                                         // scene.rs:140
        check(unsafe { tr_scene_set_light_direction(self.h, d.as_ptr()) })
    }
    pub fn set_camera(&mut self, look_from: [f32; 3], look_at: [f32; 3], up: [f32; 3]) -> Result<(), Error> {         // scene.rs:145
        check(unsafe { tr_scene_set_camera(self.h, look_from.as_ptr(), look_at.as_ptr(), up.as_ptr()) })
    }
    pub fn render(&mut self) -> Result<(), Error> { check(unsafe { tr_scene_render(self.h) }) }                      // scene.rs:151
    /// get_frame_buffer (scene.rs:92): RGB8, top-down, width*height*3 bytes (wrap in image::RgbImage::from_raw).
    pub fn get_frame_buffer(&self) -> Result<Vec<u8>, Error> {
        let mut v = vec![0u8; self.width as usize * self.height as usize * 3];
        check(unsafe { tr_scene_get_frame_buffer(self.h, v.as_mut_ptr()) })?;
        Ok(v)
    }
    pub fn get_z_buffer(&mut self) -> Result<Vec<u8>, Error> {                                                        // scene.rs:101
        let mut v = vec![0u8; self.width as usize * self.height as usize * 3];
        check(unsafe { tr_scene_get_z_buffer(self.h, v.as_mut_ptr()) })?;
        Ok(v)
    }
    pub fn get_shadow_buffer(&mut self) -> Result<Vec<u8>, Error> {                                                   // scene.rs:115
        let mut v = vec![0u8; self.width as usize * self.height as usize * 3];
        check(unsafe { tr_scene_get_shadow_buffer(self.h, v.as_mut_ptr()) })?;
        Ok(v)
    }
}

impl Drop for Scene {
    fn drop(&mut self) { unsafe { tr_scene_destroy(self.h) } }
}

// ---- the shader-trait surface ------------------------------------------------------------------------------------
// The reference's pluggable unit is `ShaderPass { prepare, vertex, fragment }` (src/scene/shader.rs:84-88): three boxed
// closures over a shared `Buffer`, assembled per pipeline name by `ShaderPipeline::new` (shader.rs:97-112).  The backend
// keeps the three roles, as compile-time roles of one C++ struct per pipeline (csrc/pipelines/*.cuh):
//
//     prepare   host (K0)    fn(&mut Uniforms, w, h, light, from, at, up) -> status       shader.rs:55-63  `Prepare`
//     vertex    device (K1)  fn(&Uniforms, &VertexIn, &mut [f32; 24] varyings)            shader.rs:66-72  `VertexShader`
//     fragment  device (K5)  fn(&ShadeArgs, &Uniforms, &FragmentIn) -> packed RGB         shader.rs:75-80  `FragmentShader`
//
// Device code cannot be supplied from Rust at run time, so on this side the surface is descriptive: the table below
// mirrors csrc/pipeline_table.cuh (same order = same ids), and `Scene::from_dir(.., name)` / `tr_scene_set_pipeline`
// select an entry by name exactly like `ShaderPipeline::new`.  Adding a pipeline means one header under csrc/pipelines/
// plus one line in TR_PIPELINES — and one line here.

/// Which host-side prepare a pass runs (shader.rs:183-279).
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Prepare { Default, ShadowPass1, ShadowPass2 }

/// One pass of a pipeline, by its three roles.
#[derive(Clone, Copy, Debug)]
pub struct ShaderPass {
    pub prepare: Prepare,
    /// whether the vertex role culls back faces (pass 1 of shadow / occlusion does not, shader.rs:671-692)
    pub vertex_culls: bool,
    /// number of f32 varyings the vertex role hands to the fragment role per polygon
    pub varying_words: u32,
    /// maps the fragment role samples: texture, normal_map, normal_map_tangent, specular_map
    pub maps: [bool; 4],
}

/// A named pipeline = its passes (shader.rs:97-112).
#[derive(Clone, Copy, Debug)]
pub struct ShaderPipeline { pub name: &'static str, pub alias: Option<&'static str>, pub passes: &'static [ShaderPass] }

const LIGHT_DEPTH: ShaderPass = ShaderPass { prepare: Prepare::ShadowPass1, vertex_culls: false, varying_words: 0, maps: [false; 4] };

/// Same order as TR_PIPELINES in csrc/pipeline_table.cuh.
pub const PIPELINES: &[ShaderPipeline] = &[
    ShaderPipeline { name: "default", alias: None, passes: &[ShaderPass { prepare: Prepare::Default, vertex_culls: true, varying_words: 3, maps: [true, false, false, false] }] },
    ShaderPipeline { name: "phong", alias: None, passes: &[ShaderPass { prepare: Prepare::Default, vertex_culls: true, varying_words: 3, maps: [true, false, false, false] }] },
    ShaderPipeline { name: "normal_map", alias: Some("true_normal"), passes: &[ShaderPass { prepare: Prepare::Default, vertex_culls: true, varying_words: 0, maps: [true, true, false, false] }] },
    ShaderPipeline { name: "specular", alias: None, passes: &[ShaderPass { prepare: Prepare::Default, vertex_culls: true, varying_words: 0, maps: [true, true, false, true] }] },
    ShaderPipeline { name: "darboux", alias: None, passes: &[ShaderPass { prepare: Prepare::Default, vertex_culls: true, varying_words: 21, maps: [true, false, true, false] }] },
    ShaderPipeline { name: "shadow", alias: None, passes: &[LIGHT_DEPTH, ShaderPass { prepare: Prepare::ShadowPass2, vertex_culls: true, varying_words: 3, maps: [true, false, false, false] }] },
    ShaderPipeline { name: "occlusion", alias: None, passes: &[LIGHT_DEPTH, ShaderPass { prepare: Prepare::ShadowPass2, vertex_culls: true, varying_words: 0, maps: [false; 4] }] },
];

/// `ShaderPipeline::new`'s name lookup (shader.rs:100-108), with the README's `true_normal` spelling.
pub fn pipeline_by_name(name: &str) -> Option<&'static ShaderPipeline> {
    PIPELINES.iter().find(|p| p.name == name || p.alias == Some(name))
}
