//! src/io.rs:28-74,152-201,264-364 of the reference: OBJ / STL loaders feeding `Mesh::new` (host-side parsing that runs
//! once; kept for API parity — the Python host's rpt_b200/io.py is the tested twin).
use std::io::{self, BufRead, BufReader, Read};

use crate::shape::{Mesh, Triangle};

fn parse_index(value: &str, len: usize) -> Option<usize> {
    // src/io.rs:11-19: 1-based, negative = relative to the end
    value.parse::<i32>().ok().and_then(|index| if index > 0 { Some((index - 1) as usize) } else { Some((len as i32 + index) as usize) })
}

/// Load a mesh from a Wavefront .OBJ file (src/io.rs:28-74): `v`, `vn`, `f` with fan triangulation; faces without
/// normals get the face normal
pub fn load_obj(reader: impl Read) -> io::Result<Mesh> {
    let mut vertices: Vec<glm::DVec3> = Vec::new();
    let mut normals: Vec<glm::DVec3> = Vec::new();
    let mut triangles = Vec::new();
    for line in BufReader::new(reader).lines() {
        let line = line?;
        let line = line.trim();
        let tokens: Vec<&str> = line.split_whitespace().collect();
        if tokens.is_empty() {
            continue;
        }
        let point = |t: &[&str]| -> io::Result<glm::DVec3> {
            let c: Vec<f64> = t[1..].iter().filter_map(|s| s.parse().ok()).collect();
            if c.len() < 3 {
                return Err(io::Error::new(io::ErrorKind::InvalidData, "could not parse OBJ point"));
            }
            Ok(glm::vec3(c[0], c[1], c[2]))
        };
        match tokens[0] {
            "v" => vertices.push(point(&tokens)?),
            "vn" => normals.push(point(&tokens)?),
            "f" => {
                let mut vi = Vec::new();
                let mut ni = Vec::new();
                for vertex in &tokens[1..] {
                    let mut args = vertex.split('/');
                    vi.push(args.next().and_then(|s| parse_index(s, vertices.len())).ok_or_else(|| io::Error::new(io::ErrorKind::InvalidData, "bad face index"))?);
                    let _texture = args.next();
                    if let Some(n) = args.next().and_then(|s| parse_index(s, normals.len())) {
                        ni.push(n);
                    }
                }
                for i in 1..(vi.len().saturating_sub(1)) {
                    let (a, b, c) = (vertices[vi[0]], vertices[vi[i]], vertices[vi[i + 1]]);
                    if ni.len() == vi.len() {
                        triangles.push(Triangle { v1: a, v2: b, v3: c, n1: normals[ni[0]], n2: normals[ni[i]], n3: normals[ni[i + 1]] });
                    } else {
                        triangles.push(Triangle::from_vertices(a, b, c));
                    }
                }
            }
            _ => {}
        }
    }
    Ok(Mesh::new(triangles))
}

/// Load a mesh from a binary .STL file (src/io.rs:264-364): 80-byte header, u32 count, 50-byte records
pub fn load_stl(mut reader: impl Read) -> io::Result<Mesh> {
    let mut data = Vec::new();
    reader.read_to_end(&mut data)?;
    if data.len() < 84 {
        return Err(io::Error::new(io::ErrorKind::InvalidData, "STL file too short"));
    }
    let count = u32::from_le_bytes([data[80], data[81], data[82], data[83]]) as usize;
    if data.len() < 84 + 50 * count {
        return Err(io::Error::new(io::ErrorKind::InvalidData, "STL file truncated"));
    }
    let f = |o: usize| f32::from_le_bytes([data[o], data[o + 1], data[o + 2], data[o + 3]]) as f64;
    let mut triangles = Vec::with_capacity(count);
    for i in 0..count {
        let base = 84 + 50 * i + 12; // skip the stored normal: the reference recomputes it from the vertices
        let p = |k: usize| glm::vec3(f(base + 12 * k), f(base + 12 * k + 4), f(base + 12 * k + 8));
        triangles.push(Triangle::from_vertices(p(0), p(1), p(2)));
    }
    Ok(Mesh::new(triangles))
}
