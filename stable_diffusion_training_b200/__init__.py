"""B200-native (sm_100a) denoising train step of lodestone-rock/stable_diffusion_training behind the reference's call
surface: train_step / TrainState (training_utils), FlaxDDPMScheduler (schedulers), lion_8bit (lion_quant).
All arithmetic runs in the in-tree CUDA library libsdb200.so (include/sdb200.h); there is no CPU or eager fallback."""
__version__ = "0.1.0"
