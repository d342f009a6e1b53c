"""Does torch's CUDA avg_pool2d backward differ on the permuted (channels-last) views the reference's loss hands it
(src/training/trainer.py:487-490) from its CPU result / its CUDA result on contiguous inputs?  (DESIGN.md §7 claim from
round 1; VERDICT r1 item 3c asks for the reproducer.)  Prints the deviations; tests/test_gpu_loss.py asserts them."""
import torch
import torch.nn.functional as F


def ssim(a, b):
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    mu1, mu2 = F.avg_pool2d(a, 3, 1, padding=1), F.avg_pool2d(b, 3, 1, padding=1)
    s1 = F.avg_pool2d(a * a, 3, 1, padding=1) - mu1.pow(2)
    s2 = F.avg_pool2d(b * b, 3, 1, padding=1) - mu2.pow(2)
    s12 = F.avg_pool2d(a * b, 3, 1, padding=1) - mu1 * mu2
    return (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))).mean()


def grad(x, gt, dev, contiguous, dtype):
    x = x.to(dev, dtype).requires_grad_(True)
    g = gt.to(dev, dtype)
    a, b = x.permute(2, 0, 1).unsqueeze(0), g.permute(2, 0, 1).unsqueeze(0)
    if contiguous:
        a, b = a.contiguous(), b.contiguous()
    val = ssim(a, b)
    val.backward()
    return val.item(), x.grad.detach().cpu().double()


if __name__ == "__main__":
    print("torch", torch.__version__)
    for H, W in ((16, 32), (90, 160), (1080, 1920)):
        gen = torch.Generator().manual_seed(H)
        x, gt = torch.rand(H, W, 3, generator=gen), torch.rand(H, W, 3, generator=gen)
        v_cpu, g_cpu = grad(x, gt, "cpu", False, torch.float64)
        for dev in (["cuda"] if torch.cuda.is_available() else []):
            for contiguous in (False, True):
                for dt in (torch.float32, torch.float64):
                    v, g = grad(x, gt, dev, contiguous, dt)
                    dev_max = (g - g_cpu).abs().max().item() / g_cpu.abs().max().item()
                    l2 = ((g - g_cpu).norm() / g_cpu.norm()).item()
                    print(f"{H}x{W} {dev} contiguous={contiguous} {dt}: value diff {abs(v - v_cpu):.2e}; grad max-dev / max {dev_max:.3e}; L2-rel {l2:.3e}")
        v, g = grad(x, gt, "cpu", True, torch.float64)
        print(f"{H}x{W} cpu contiguous=True: grad L2-rel vs permuted cpu {((g - g_cpu).norm() / g_cpu.norm()).item():.2e}")
