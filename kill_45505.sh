ssh test@127.0.0.1 "kill -9 5696"
ssh test@127.0.0.1 "kill -9 5697"
kill -9 5044
