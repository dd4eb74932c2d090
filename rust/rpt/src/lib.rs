//! `rpt` — the reference's path tracer API (neevparikh/rpt `src/lib.rs:6-20`) over the B200 CUDA core.
//!
//! Same names, argument meaning and panics as the reference; the Monte Carlo light transport itself runs in
//! `librptx.so` (`include/rptx.h`).  See README.md: this crate is source only in this repository.
#![allow(clippy::too_many_arguments)]

pub use buffer::*;
pub use camera::*;
pub use color::*;
pub use environment::*;
pub use io::*;
pub use kdtree::*;
pub use light::*;
pub use material::*;
pub use medium::*;
pub use object::*;
pub use renderer::*;
pub use scene::*;
pub use shape::*;
pub use {glm, image};

mod buffer;
mod camera;
mod color;
mod environment;
mod ffi;
mod io;
mod kdtree;
mod light;
mod lower;
mod material;
mod medium;
mod object;
mod renderer;
mod scene;
mod shape;
