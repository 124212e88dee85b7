"""small invocations of every round-2 kernel, for compute-sanitizer (memcheck):
  compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import sys
import torch
sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe

g = torch.Generator().manual_seed(0)
x = (0.2 * torch.randn(3, 16000 * 3 + 4, generator=g)).clamp(-1, 1)
lens = [x.shape[1], 20001, 777]
o = []
o.append(fe.whisper_log_mel_batch(x.cuda(), n_mels=128))
o.append(fe.whisper_log_mel_batch(x.cuda(), lengths=lens, n_mels=80, padding=48000, out_dtype=torch.float16))
o.append(fe.whisper_log_mel_batch(x.cuda(), lengths=lens, n_mels=128, padding=480000))
o.append(fe.whisper_log_mel_windows(x.cuda(), lengths=lens, n_mels=80))
o.append(fe.whisper_log_mel_batch(x.cuda()[:, 1:], n_mels=80))                      # unaligned rows: frame-pair kernel
o.append(fe.kaldi_fbank_batch(x.cuda(), lengths=lens))
o.append(fe.cosy_mel_batch(x.cuda(), lengths=[x.shape[1], 20001, 1000]))
o.append(fe.wav2lip_mel_batch(x.cuda(), lengths=lens))
o.append(fe.wav2lip_mel_batch(x.cuda(), pad_mode="constant", frame_major=True))
kw = dict(fe.REFERENCE_RESAMPLER)
y = (0.2 * torch.randn(2, 44100 + 4, generator=g)).clamp(-1, 1)
o.append(fe.condition_pcm(y.cuda(), 44100, 16000, lengths=[44104, 9000], **kw))     # tensor-core kernel, float32 mono
p16 = torch.randint(-20000, 20000, (2, 44104, 2), generator=g, dtype=torch.int16)
o.append(fe.condition_pcm(p16.cuda(), 44100, 16000, **kw))                           # int16 stereo frames read by the producers
o.append(fe.condition_pcm(p16.cuda()[:, :44101].contiguous(), 44100, 16000, **kw))   # unaligned rows: mono pre-pass
o.append(fe.condition_pcm(y.cuda(), 48000, 16000, **kw))                             # regrouped blocks
o.append(fe.condition_pcm(p16.cuda(), 44100, 16000))                                 # CUDA-core kernel
o.append(fe.condition_stereo_auto(p16.cuda(), 44100, 16000, lengths=[44104, 30000]))
o.append(fe.pcm_whisper_log_mel_batch(p16.cuda(), 44100, lengths=[44104, 30000], n_mels=128, padding=4800))
o.append(fe.pcm_whisper_log_mel_host(p16, 44100, n_mels=80))
torch.cuda.synchronize()
print("ok", sum(float(t.float().abs().sum()) for t in o))
