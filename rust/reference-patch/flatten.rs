//! Goes to `src/scene/triangle_bvh/flatten.rs` of bluecube/minipath (`mod flatten; pub use flatten::FlatBvh;` in
//! `triangle_bvh/mod.rs`) — inside the module because `TriangleBvh`'s fields and `InnerNode` are private (mod.rs:20-38).
//! Hands the BVH to libminipath_b200 in the layout of `MpFlatBvh`: the node and packet arrays as they are (96 B of u16x8
//! bounds + 8 links; 9 x u16x8), plus the enclosing box every node / packet is dequantised against, which the reference
//! carries on its traversal stack (ray_bvh_intersection.rs:21-23) and the device reads from a table instead.
//! NOT compiled in the minipath_b200 repository (no Rust toolchain there).
use minipath_b200_sys as sys;

use super::{CompressedNodeLink, NodeLink, TriangleBvh};

pub struct FlatBvh {
    pub raw: sys::MpFlatBvh, // points into the vectors below
    _inner: Vec<sys::MpInnerNode>,
    _inner_box: Vec<sys::MpBox>,
    _packets: Vec<sys::MpTrianglePacket>,
    _packet_box: Vec<sys::MpBox>,
    _shading: Vec<sys::MpTriShading>,
    _vertices: Vec<sys::MpVertexShading>,
}
// the raw pointers refer to heap memory owned by the same value
unsafe impl Send for FlatBvh {}
unsafe impl Sync for FlatBvh {}

const K: f32 = 1.0 / 65535.0; // compressed_geometry.rs:49

/// One lane of `RelativePoint8::decompress` (compressed_geometry.rs:95-110): `size.mul_add(q as f32 * K, min)`.
#[inline]
fn dequantise(q: u16, min: f32, size: f32) -> f32 {
    size.mul_add(q as f32 * K, min)
}

impl TriangleBvh {
    pub fn flatten(&self) -> FlatBvh {
        let inner: Vec<sys::MpInnerNode> = self
            .inner_nodes
            .iter()
            .map(|n| {
                let b = &n.child_bounds;
                sys::MpInnerNode {
                    bounds: [
                        b.min.x.0.to_array(), b.min.y.0.to_array(), b.min.z.0.to_array(),
                        b.max.x.0.to_array(), b.max.y.0.to_array(), b.max.z.0.to_array(),
                    ],
                    links: n.child_links.map(|l| l.0),
                }
            })
            .collect();
        let packets: Vec<sys::MpTrianglePacket> = self
            .triangle_geometry
            .iter()
            .map(|t| sys::MpTrianglePacket {
                v: [0, 1, 2].map(|k| [t.0[k].x.0.to_array(), t.0[k].y.0.to_array(), t.0[k].z.0.to_array()]),
            })
            .collect();
        let mut inner_box = vec![sys::MpBox::default(); inner.len()];
        let mut packet_box = vec![sys::MpBox::default(); packets.len()];

        // walk from (root, bounding_box) exactly like TriangleBvh::intersect pushes children (ray_bvh_intersection.rs:34-37,149-162)
        let bb = &self.bounding_box;
        let root_box = sys::MpBox {
            min: [bb.min.x, bb.min.y, bb.min.z],
            _pad0: 0.0,
            size: [bb.max.x - bb.min.x, bb.max.y - bb.min.y, bb.max.z - bb.min.z],
            _pad1: 0.0,
        };
        let mut stack = vec![(self.root, root_box)];
        while let Some((link, enclosing)) = stack.pop() {
            match link.decode() {
                NodeLink::Null => {}
                NodeLink::Leaf { indices } => {
                    for p in indices {
                        packet_box[usize::from(p)] = enclosing;
                    }
                }
                NodeLink::Inner { index } => {
                    let i = usize::from(index);
                    inner_box[i] = enclosing;
                    let node = &inner[i];
                    for lane in 0..8 {
                        let child = CompressedNodeLink(node.links[lane]);
                        if child == CompressedNodeLink::NULL {
                            continue;
                        }
                        let mut b = sys::MpBox::default();
                        for a in 0..3 {
                            let lo = dequantise(node.bounds[a][lane], enclosing.min[a], enclosing.size[a]);
                            let hi = dequantise(node.bounds[3 + a][lane], enclosing.min[a], enclosing.size[a]);
                            b.min[a] = lo;
                            b.size[a] = hi - lo; // AABBSized::from(&AABB), aabb.rs:292-303
                        }
                        stack.push((child, b));
                    }
                }
            }
        }

        let shading: Vec<sys::MpTriShading> = self
            .triangle_shading_data
            .iter()
            .map(|s| sys::MpTriShading {
                v: [0, 1, 2].map(|k| u32::try_from(s.vertex_indices.0[k]).expect("vertex index exceeds u32")),
                flags_material: (u32::try_from(s.material).expect("material exceeds u32") & !sys::MP_TRI_FLAT)
                    | if s.flat_shading { sys::MP_TRI_FLAT } else { 0 },
            })
            .collect();
        let vertices: Vec<sys::MpVertexShading> = self
            .vertex_data
            .iter()
            .map(|v| sys::MpVertexShading {
                normal: [v.normal.x, v.normal.y, v.normal.z],
                _pad0: 0.0,
                tex: [v.texture_coords.x, v.texture_coords.y, v.texture_coords.z],
                _pad1: 0.0,
            })
            .collect();
        let n_triangles = self
            .triangle_geometry
            .iter()
            .map(|t| 8 - t.0[0].is_zero().bitmask().count_ones().min(8)) // padding lanes are all-zero (building.rs:157)
            .sum::<u32>();
        let raw = sys::MpFlatBvh {
            bbox_min: [bb.min.x, bb.min.y, bb.min.z],
            bbox_max: [bb.max.x, bb.max.y, bb.max.z],
            root: self.root.0,
            n_inner: inner.len() as u32,
            n_packets: packets.len() as u32,
            n_vertices: vertices.len() as u32,
            n_triangles,
            n_materials: 0, // the reference writes material: 0 everywhere (building.rs:201): no table, constant grey
            inner: inner.as_ptr(),
            inner_box: inner_box.as_ptr(),
            packets: packets.as_ptr(),
            packet_box: packet_box.as_ptr(),
            tri_shading: shading.as_ptr(),
            vertices: vertices.as_ptr(),
            materials: std::ptr::null(),
        };
        FlatBvh { raw, _inner: inner, _inner_box: inner_box, _packets: packets, _packet_box: packet_box, _shading: shading, _vertices: vertices }
    }
}
