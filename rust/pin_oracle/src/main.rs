//! pin_oracle — the reference's `Scene` (src/scene.rs, src/scene/shader.rs, src/scene/util.rs, included UNCHANGED by
//! path) driven headless through the golden cases of this repository, printing one JSON object per case with the
//! sha256 of what its public getters return:
//!
//!   frame_topdown  Scene::get_frame_buffer()   scene.rs:92-97    (RGB8, top-down)
//!   z_gray         Scene::get_z_buffer()       scene.rs:101-111  (`f32 as u8` gray view of the depth buffer)
//!   shadow_gray    Scene::get_shadow_buffer()  scene.rs:115-125  (two-pass pipelines)
//!
//! Per case the sequence is the app loop's (app.rs:170, 208-210): clear, set_light_direction, set_camera, render.
//! Camera and light vectors are read as f32 BIT PATTERNS from cases.txt (written by tests/golden/make_golden.py), so no
//! sin/cos of this platform enters.  Usage, from a checkout of the reference with this directory copied to
//! <reference>/pin_oracle:
//!
//!   cd <reference>/pin_oracle && cp ../Cargo.lock . && cargo run --release -- ../assets cases.txt > pinned.json
//!   python <this repo>/tools/pin_diff.py pinned.json
//!
//! UNBUILT ON THIS BOX (no cargo/rustc in the build image).
#[path = "../../src/scene.rs"]
mod scene;

use std::fs::File;
use std::io::{BufRead, BufReader};

use nalgebra::Vector3;
use obj::raw::parse_obj;
use scene::Scene;

// ---- SHA-256 (FIPS 180-4), so that no crate beyond the reference's own is needed ---------------------------------
const K: [u32; 64] = [
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01,
    0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc,
    0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147,
    0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08,
    0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208,
    0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2,
];

fn sha256_hex(data: &[u8]) -> String {
    let mut h: [u32; 8] = [0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19];
    let mut msg = data.to_vec();
    let bits = (data.len() as u64).wrapping_mul(8);
    msg.push(0x80);
    while msg.len() % 64 != 56 {
        msg.push(0);
    }
    msg.extend_from_slice(&bits.to_be_bytes());
    let mut w = [0u32; 64];
    for block in msg.chunks(64) {
        for i in 0..16 {
            w[i] = u32::from_be_bytes([block[4 * i], block[4 * i + 1], block[4 * i + 2], block[4 * i + 3]]);
        }
        for i in 16..64 {
            let s0 = w[i - 15].rotate_right(7) ^ w[i - 15].rotate_right(18) ^ (w[i - 15] >> 3);
            let s1 = w[i - 2].rotate_right(17) ^ w[i - 2].rotate_right(19) ^ (w[i - 2] >> 10);
            w[i] = w[i - 16].wrapping_add(s0).wrapping_add(w[i - 7]).wrapping_add(s1);
        }
        let mut v = h;
        for i in 0..64 {
            let s1 = v[4].rotate_right(6) ^ v[4].rotate_right(11) ^ v[4].rotate_right(25);
            let ch = (v[4] & v[5]) ^ (!v[4] & v[6]);
            let t1 = v[7].wrapping_add(s1).wrapping_add(ch).wrapping_add(K[i]).wrapping_add(w[i]);
            let s0 = v[0].rotate_right(2) ^ v[0].rotate_right(13) ^ v[0].rotate_right(22);
            let maj = (v[0] & v[1]) ^ (v[0] & v[2]) ^ (v[1] & v[2]);
            let t2 = s0.wrapping_add(maj);
            v = [t1.wrapping_add(t2), v[0], v[1], v[2], v[3].wrapping_add(t1), v[4], v[5], v[6]];
        }
        for i in 0..8 {
            h[i] = h[i].wrapping_add(v[i]);
        }
    }
    h.iter().map(|x| format!("{:08x}", x)).collect()
}

fn f32_from_hex(s: &str) -> f32 {
    f32::from_bits(u32::from_str_radix(s, 16).expect("cases.txt: bad hex word"))
}

fn main() -> Result<(), Box<dyn std::error::Error>> {
    let args: Vec<String> = std::env::args().collect();
    if args.len() != 3 {
        eprintln!("usage: pin_oracle <assets dir> <cases.txt>");
        std::process::exit(2);
    }
    assert_eq!(sha256_hex(b"abc"), "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad");
    let mut first = true;
    println!("[");
    for line in BufReader::new(File::open(&args[2])?).lines() {
        let line = line?;
        let t: Vec<&str> = line.split_whitespace().collect();
        if t.is_empty() || t[0].starts_with('#') {
            continue;
        }
        // asset pipeline width height  look_from(3) look_at(3) up(3) light(3) as f32 bit patterns
        assert_eq!(t.len(), 16, "cases.txt: 16 fields per line");
        let (asset, pipeline) = (t[0], t[1]);
        let (w, h): (u32, u32) = (t[2].parse()?, t[3].parse()?);
        let v: Vec<f32> = t[4..16].iter().map(|s| f32_from_hex(s)).collect();
        let dir = format!("{}/{}", args[1], asset);
        // app::run's load sequence, app.rs:87-131
        let obj = parse_obj(BufReader::new(File::open(format!("{}/model.obj", dir))?))?;
        let texture = image::open(format!("{}/texture.tga", dir))?.into_rgb8();
        let normal_map = image::open(format!("{}/normal_map.tga", dir))?.into_rgb8();
        let normal_map_tangent = image::open(format!("{}/normal_map_tangent.tga", dir))?.into_rgb8();
        let specular_map = image::open(format!("{}/specular_map.tga", dir))?.into_rgb8();
        let mut scene = Scene::new(w, h, obj, texture, normal_map, normal_map_tangent, specular_map, pipeline.to_string());
        scene.clear();
        scene.set_light_direction(Vector3::new(v[9], v[10], v[11]));
        scene.set_camera(Vector3::new(v[0], v[1], v[2]), Vector3::new(v[3], v[4], v[5]), Vector3::new(v[6], v[7], v[8]));
        scene.render();
        let frame = sha256_hex(scene.get_frame_buffer().as_raw());
        let z = sha256_hex(scene.get_z_buffer().as_raw());
        let shadow = sha256_hex(scene.get_shadow_buffer().as_raw());
        if !first {
            println!(",");
        }
        first = false;
        print!(
            " {{\"case\": [\"{}\", \"{}\", {}, {}], \"frame_topdown\": \"{}\", \"z_gray\": \"{}\", \"shadow_gray\": \"{}\"}}",
            asset, pipeline, w, h, frame, z, shadow
        );
    }
    println!("\n]");
    Ok(())
}
