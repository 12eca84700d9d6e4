ssh test@127.0.0.1 "kill -9 9475"
ssh test@127.0.0.1 "kill -9 9476"
kill -9 8873
