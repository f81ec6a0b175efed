// The reference's own integration tests with the renderer swapped (rayca-soft/tests/gltf.rs:185-214 `boxes`,
// :296-317 `flight`): `SoftRenderer` -> `CudaRenderer` is the only change.  UNCOMPILED in the authoring container.
use std::error::Error;

use rayca_cuda::CudaRenderer;
use rayca_soft::*;

#[test]
fn boxes() -> Result<(), Box<dyn Error>> {
    let mut scene = Scene::default();
    let assets = Assets::new();
    scene.push_gltf_from_path(tests::get_model_path().join("box/box.gltf"), &assets)?;
    scene.push_model(CudaRenderer::create_default_model());

    let mut image = Image::new(128, 128, ColorType::RGBA8);
    let mut soft_image = Image::new(128, 128, ColorType::RGBA8);
    let config = Config::builder().max_depth(1).build();
    CudaRenderer::new_with_config(config).draw(&scene, &mut image);
    SoftRenderer::new_with_config(config).draw(&scene, &mut soft_image);
    // RGB within one 8-bit step (1e-3 absolute before quantisation)
    for (a, b) in image.bytes().iter().zip(soft_image.bytes()) {
        assert!((*a as i32 - *b as i32).abs() <= 1);
    }
    image.dump_png(tests::get_artifact_path().join("box-cuda.png"))?;
    Ok(())
}
