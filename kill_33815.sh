ssh test@127.0.0.1 "kill -9 5572"
ssh test@127.0.0.1 "kill -9 5573"
kill -9 5044
