#!/bin/sh
# Applies the hook-in edits flatten.rs needs to a checkout of zhouhang95/ray_tracing_in_one_weekend (run from its
# root).  UNTESTED against a Rust toolchain.  Every edit first checks that the line it is about to touch still reads
# what it read in the surveyed snapshot, so a drifted checkout fails loudly instead of being mangled.
set -eu
expect() { # file line needle
    sed -n "${2}p" "$1" | grep -q -- "$3" || { echo "$1:$2 does not contain '$3' — checkout differs from the surveyed snapshot" >&2; exit 1; }
}
pubcrate() { # file first last : prefix the field declarations on those lines with pub(crate)
    sed -i "${2},${3}s/^\(  *\)\([a-z_][a-z_0-9]*:\)/\1pub(crate) \2/" "$1"
}
# (1) supertraits: `dyn Hitable` / `dyn Material` / `dyn Texture` become flattenable without a downcast
expect src/hitable.rs 50 'pub trait Hitable: Send + Sync'
sed -i '50s/Send + Sync/Send + Sync + crate::flatten::FlattenHitable/' src/hitable.rs
expect src/material.rs 10 'pub trait Material: Send + Sync'
sed -i '10s/Send + Sync/Send + Sync + crate::flatten::FlattenMaterial/' src/material.rs
expect src/texture.rs 11 'pub trait Texture: Send + Sync'
sed -i '11s/Send + Sync/Send + Sync + crate::flatten::FlattenTexture/' src/texture.rs
# (2) private fields the flattener reads -> pub(crate)
expect src/hitable.rs 158 'pub struct BvhNode';        pubcrate src/hitable.rs 159 161
expect src/hitable.rs 438 'pub struct RotateY';        pubcrate src/hitable.rs 439 444
expect src/hitable.rs 523 'pub struct ConstantMedium'; pubcrate src/hitable.rs 524 526
expect src/texture.rs 25 'pub struct CheckerTex';      pubcrate src/texture.rs 26 27
expect src/texture.rs 53 'struct Perlin';              sed -i '53s/^struct Perlin/pub(crate) struct Perlin/' src/texture.rs
pubcrate src/texture.rs 54 57
expect src/texture.rs 149 'pub struct PerlinTex';      pubcrate src/texture.rs 150 151
expect src/texture.rs 170 'pub struct ImageTex';       pubcrate src/texture.rs 171 171
expect src/camera.rs 6 'pub struct Camera';            pubcrate src/camera.rs 7 10
# (3) the module, the dependency, and the call site in main()
cp "$(dirname "$0")/flatten.rs" src/flatten.rs
grep -q '^mod flatten;' src/main.rs || sed -i '0,/^mod /s//mod flatten;\nmod /' src/main.rs
grep -q '^rtgpu-sys' Cargo.toml || sed -i 's/^\[dependencies\]$/[dependencies]\nrtgpu-sys = { path = "..\/rtgpu-sys" }/' Cargo.toml
echo "patched: now replace the body of main() between 'let (world, cam) = ...' and the final img.save with the call"
echo "site shown in INTEGRATION.md section 5 (FlatScene::from_world -> GpuScene::new -> render -> resolve_rgb8)."
